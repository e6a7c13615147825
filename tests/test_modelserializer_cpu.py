"""coefficients.bin / updaterState.bin in ModelSerializer's zip layout (util/ModelSerializer.java:39-40, 120-163) for the flat
vectors of the BASELINE models: byte-level layout of the restated Nd4j.write format and a lossless round trip."""
import io
import json
import struct

import numpy as np


def test_nd4j_array_layout_is_big_endian_shapeinfo_then_data():
    from deeplearning4j_b200 import modelserializer as ms
    buf = io.BytesIO()
    ms.write_array(buf, np.array([1.5, -2.0, 3.25], np.float32))
    raw = buf.getvalue()
    # UTF("DIRECT") int32(8) UTF("INT") then rank=2, shape 1x3, strides 3,1, offset 0, ews 1, order 'c'
    assert raw[:2] == struct.pack(">H", 6) and raw[2:8] == b"DIRECT"
    assert struct.unpack(">i", raw[8:12]) == (8,) and raw[12:17] == struct.pack(">H", 3) + b"INT"
    assert struct.unpack(">8i", raw[17:49]) == (2, 1, 3, 3, 1, 0, 1, ord("c"))
    assert raw[49:57] == struct.pack(">H", 6) + b"DIRECT" and struct.unpack(">i", raw[57:61]) == (3,)
    assert raw[61:68] == struct.pack(">H", 5) + b"FLOAT"
    assert struct.unpack(">3f", raw[68:80]) == (1.5, -2.0, 3.25) and len(raw) == 80
    back = ms.read_array(io.BytesIO(raw))
    assert back.shape == (1, 3) and back.tolist() == [[1.5, -2.0, 3.25]]


def test_model_zip_round_trip_of_the_flat_vectors(tmp_path, oracle):
    from deeplearning4j_b200 import modelserializer as ms, zoo
    m = zoo.lenet()
    P, layout = zoo.param_layout(m, oracle.out_size)
    params = zoo.init_params(m, out_size=oracle.out_size)
    state = np.random.default_rng(0).random(2 * P).astype(np.float32)  # AdaDelta: msg | msdx per block
    conf = json.dumps({"model": m["name"], "P": P, "layout": [[n, k, off, list(shp)] for n, k, off, shp, _ in layout]})
    path = tmp_path / "lenet.zip"
    ms.write_model(path, conf, params, state)
    conf2, p2, s2 = ms.restore_model(path)
    assert conf2 == conf and p2.dtype == np.float32
    assert np.array_equal(p2, params) and np.array_equal(s2, state)
    import zipfile
    assert set(zipfile.ZipFile(path).namelist()) == {"configuration.json", "coefficients.bin", "updaterState.bin"}
    # a block read back through the layout is the block that was written (conv: bias | W[c_out,c_in,kH,kW] 'c')
    for n, k, off, shp, _ in layout:
        cnt = int(np.prod(shp))
        assert np.array_equal(p2[off:off + cnt].reshape(shp), params[off:off + cnt].reshape(shp))
