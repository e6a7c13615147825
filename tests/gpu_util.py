"""Shared helpers for the -m gpu parity tests (call the C ABI through deeplearning4j_b200._lib)."""
import ctypes as C

import numpy as np


def gpu_available():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def rel_err(got, ref):
    """max |got-ref| / max(|ref|, tiny): the 'relative to max|ref| per tensor' metric of BASELINE.md §4."""
    got, ref = np.asarray(got, np.float64), np.asarray(ref, np.float64)
    return float(np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))


def to_nhwc(x, cs=None):
    """NCHW numpy -> NHWC numpy with pixel stride cs (zero padded channels)."""
    B, Cc, H, W = x.shape
    cs = cs or (Cc + 3) // 4 * 4
    out = np.zeros((B, H, W, cs), np.float32)
    out[..., :Cc] = x.transpose(0, 2, 3, 1)
    return out


def from_nhwc(a, Cc):
    return np.ascontiguousarray(a[..., :Cc].transpose(0, 3, 1, 2))


class Dev:
    """Tiny device-buffer helper over dl4jb200_malloc / memcpy."""

    def __init__(self, ctx):
        self.ctx = ctx
        self.ptrs = []

    def put(self, arr):
        arr = np.ascontiguousarray(arr)
        p = self.ctx.malloc(max(arr.nbytes, 16))
        self.ptrs.append(p)
        from deeplearning4j_b200 import _lib
        _lib.call("dl4jb200_memcpy_h2d", self.ctx.handle, p, arr.ctypes.data_as(C.c_void_p), arr.nbytes)
        return p

    def empty(self, nbytes, fill=None):
        p = self.ctx.malloc(max(nbytes, 16))
        self.ptrs.append(p)
        if fill is not None:
            from deeplearning4j_b200 import _lib
            _lib.call("dl4jb200_memset", self.ctx.handle, p, fill, nbytes)
        return p

    def get(self, p, shape, dtype=np.float32):
        return self.ctx.download(p, shape, dtype)

    def free(self):
        for p in self.ptrs:
            self.ctx.free(p)
        self.ptrs = []
