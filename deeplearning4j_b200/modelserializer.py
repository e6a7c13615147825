"""ModelSerializer-compatible checkpoints of the flat vectors (SURVEY.md §8 f4, §5 "checkpoint / resume").

DL4J's ModelSerializer writes a zip with `configuration.json`, `coefficients.bin` (the flat [1,P] parameter vector, in the
layout of zoo.param_layout == MultiLayerNetwork's flat view) and `updaterState.bin` (util/ModelSerializer.java:39-40,
120-163); the two .bin entries are `Nd4j.write(array, DataOutputStream)`. That serialisation lives in the un-vendored nd4j
(nd4j-api 1.0.0-SNAPSHOT, `Nd4j.write` / `BaseDataBuffer.write`), so its published layout is restated here -- parity
UNPINNED, no fixture in the reference tree: all values big-endian (java.io.DataOutputStream);
  shape-information buffer:  UTF "allocation mode", int32 length = 2*rank+4, UTF "INT",
                             int32[length] = rank, shape[rank], stride[rank], offset, elementWiseStride, order ('c' = 99)
  data buffer:               UTF "allocation mode", int32 length, UTF "FLOAT", float32[length]
(writeUTF = uint16 byte count + modified UTF-8). The device-resident step driver syncs its parameter / updater-state
vectors to the host on demand and writes them with this module, so a model trained through the shim restores in stock DL4J
(restoreMultiLayerNetwork reads the same three entries, ModelSerializer.java:207-253).
"""
import io
import struct
import zipfile

import numpy as np

ALLOCATION_MODE = "DIRECT"  # AllocationMode.DIRECT of nd4j-native; the reader accepts any


def _utf(s):
    b = s.encode("utf-8")
    return struct.pack(">H", len(b)) + b


def _read_utf(f):
    (n,) = struct.unpack(">H", f.read(2))
    return f.read(n).decode("utf-8")


def write_array(f, arr, order="c"):
    """Nd4j.write for a float32 array (row vector [1,P] for the flat parameter / updater-state vectors)."""
    a = np.ascontiguousarray(arr, dtype=np.float32)
    if a.ndim == 1:
        a = a.reshape(1, -1)
    rank = a.ndim
    strides = [int(s // 4) for s in a.strides]
    info = [rank] + list(a.shape) + strides + [0, 1, ord(order)]
    f.write(_utf(ALLOCATION_MODE) + struct.pack(">i", len(info)) + _utf("INT"))
    f.write(struct.pack(">%di" % len(info), *info))
    f.write(_utf(ALLOCATION_MODE) + struct.pack(">i", a.size) + _utf("FLOAT"))
    f.write(a.astype(">f4").tobytes())


def read_array(f):
    _read_utf(f)
    (n,) = struct.unpack(">i", f.read(4))
    kind = _read_utf(f)
    if kind != "INT":
        raise ValueError("shape information buffer expected, got " + kind)
    info = struct.unpack(">%di" % n, f.read(4 * n))
    rank = info[0]
    shape, order = info[1:1 + rank], chr(info[2 * rank + 3])
    _read_utf(f)
    (length,) = struct.unpack(">i", f.read(4))
    kind = _read_utf(f)
    if kind not in ("FLOAT", "DOUBLE"):
        raise ValueError("unsupported data buffer type " + kind)
    dt = ">f4" if kind == "FLOAT" else ">f8"
    data = np.frombuffer(f.read(length * int(dt[2])), dtype=dt).astype(np.float32)
    return data.reshape(shape, order="C" if order == "c" else "F")


def write_model(path, configuration_json, params, updater_state=None):
    """ModelSerializer.writeModel(model, file, saveUpdater)."""
    with zipfile.ZipFile(path, "w", zipfile.ZIP_DEFLATED) as z:
        z.writestr("configuration.json", configuration_json)
        buf = io.BytesIO()
        write_array(buf, params)
        z.writestr("coefficients.bin", buf.getvalue())
        if updater_state is not None and np.size(updater_state) > 0:
            buf = io.BytesIO()
            write_array(buf, updater_state)
            z.writestr("updaterState.bin", buf.getvalue())


def restore_model(path):
    """-> (configuration_json, params [P], updater_state [S] or None)."""
    with zipfile.ZipFile(path) as z:
        conf = z.read("configuration.json").decode("utf-8")
        params = read_array(io.BytesIO(z.read("coefficients.bin"))).reshape(-1)
        state = read_array(io.BytesIO(z.read("updaterState.bin"))).reshape(-1) if "updaterState.bin" in z.namelist() else None
    return conf, params, state
