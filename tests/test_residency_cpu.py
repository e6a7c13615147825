"""CPU check of the host side of the residency protocol (helpers.ResidentArray / Context.ain / aout) against a stand-in
for the registry entry points that implements the documented state machine of include/dl4jb200.h (HOST_NEWER / SYNCED /
DEVICE_NEWER): which calls the helpers make, and that nothing is downloaded until host code reads."""
import ctypes as C

import numpy as np


class _FakeRegistry:
    def __init__(self):
        self.state, self.h2d, self.d2h, self.log = {}, 0, 0, []

    def call(self, name, *a):
        self.log.append(name)
        key = a[1].value if len(a) > 1 and isinstance(a[1], C.c_void_p) else None
        if name == "dl4jb200_buf_acquire":
            _, _, nbytes, access, out = a
            st = self.state.setdefault(key, 0)
            if access & 1 and st == 0:
                self.h2d += nbytes
                self.state[key] = 1
            if access & 2:
                self.state[key] = 2
            out._obj.value = 0x1000 + len(self.state)
        elif name == "dl4jb200_buf_register":
            self.state.setdefault(key, 0)
            a[3]._obj.value = 0x1000 + len(self.state)
        elif name == "dl4jb200_buf_mark_host_dirty":
            self.state[key] = 0
        elif name == "dl4jb200_buf_sync_to_host":
            if self.state[key] == 2:
                self.d2h += 1
                self.state[key] = 1
        elif name == "dl4jb200_buf_state":
            a[2]._obj.value = self.state.get(key, -1)
        elif name == "dl4jb200_buf_release":
            self.state.pop(key, None)


def test_resident_array_protocol(monkeypatch):
    from deeplearning4j_b200 import _lib, helpers as H
    fake = _FakeRegistry()
    monkeypatch.setattr(_lib, "call", fake.call)

    class Ctx:
        handle = C.c_void_p(1)
        resident = True
    ctx = Ctx()
    r = H.ResidentArray(ctx, (2, 3, 4, 5))
    assert r.state() == _lib.BUF_DEVICE_NEWER and r.shape == (2, 3, 4, 5) and r.ndim == 4 and r.size == 120
    assert fake.d2h == 0
    r.device_ptr(_lib.BUF_READ)                    # handed to another helper: no traffic
    assert (fake.h2d, fake.d2h) == (0, 0)
    _ = r + 1.0                                    # host arithmetic: one download ...
    assert fake.d2h == 1 and r.state() == _lib.BUF_SYNCED
    _ = r.max(), np.asarray(r), r[0, 0]            # ... and only one
    assert fake.d2h == 1
    r.numpy()[...] = 3.0
    r.mark_host_dirty()
    r.device_ptr(_lib.BUF_READ)                    # the next kernel that reads it uploads
    assert fake.h2d == max(r.nbytes, 16) and r.state() == _lib.BUF_SYNCED
    r.device_ptr(_lib.BUF_READ | _lib.BUF_WRITE)   # in-place kernel
    assert r.state() == _lib.BUF_DEVICE_NEWER and fake.h2d == max(r.nbytes, 16)
    x = np.arange(24, dtype=np.float32).reshape(2, 3, 4)
    xr = H.ResidentArray.from_host(ctx, x)
    assert xr.state() == _lib.BUF_HOST_NEWER and xr.numpy() is not None and fake.d2h == 1
    key = xr._key.value
    del xr
    assert key not in fake.state                   # released when the array dies
