"""-m gpu: backprop-data of strided convolutions with kernels larger than 1x1 -- decomposed by stride phase onto the TMA /
tcgen05 path (conv.cu bwd_data_phases) -- against the oracle, with odd / even sizes, padding, SAME mode and mixed strides."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import Dev, from_nhwc, rel_err, to_nhwc

pytestmark = pytest.mark.gpu

CASES = [  # B, Cin, H, W, Cout, k, s, pad, mode
    (4, 64, 28, 28, 192, (5, 5), (2, 2), (2, 2), "Truncate"),   # AlexNet cnn2
    (3, 16, 15, 17, 64, (5, 5), (2, 2), (2, 2), "Truncate"),    # odd sizes
    (2, 32, 14, 14, 32, (3, 3), (2, 2), (0, 0), "Same"),        # SAME, asymmetric padding
    (2, 8, 13, 13, 96, (3, 3), (2, 2), (1, 1), "Truncate"),
    (2, 24, 12, 16, 64, (4, 4), (2, 2), (1, 1), "Truncate"),    # even kernel
    (2, 12, 17, 17, 32, (3, 3), (3, 3), (0, 0), "Truncate"),    # stride == kernel
    (2, 20, 16, 11, 64, (3, 5), (2, 1), (1, 2), "Truncate"),    # stride along H only
    (2, 16, 32, 32, 32, (7, 7), (4, 4), (3, 3), "Truncate"),    # stride 4: sixteen phases
    (2, 16, 9, 9, 32, (2, 2), (3, 3), (0, 0), "Truncate"),      # kernel smaller than the stride: gather fallback
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("layout_name", ["NHWC", "NCHW"])
def test_strided_backprop_data_matches_oracle(oracle, case, layout_name):
    from deeplearning4j_b200 import _lib as L, helpers as H
    o = oracle
    B, Cin, Hh, W, Cout, k, s, pad, mode = case
    rng = np.random.default_rng(23)
    x = rng.standard_normal((B, Cin, Hh, W)).astype(np.float32)
    w = (rng.standard_normal((Cout, Cin) + k) * 0.2).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, k, s, pad, (1, 1), mode)
    delta = rng.standard_normal((B, Cout, d.out_h, d.out_w)).astype(np.float32)
    omode = {"Same": o.SAME, "Truncate": o.TRUNCATE}[mode]
    dwr, dbr, dxr = o.conv_bwd(x, w, delta, stride=s, pad=pad, mode=omode)
    nhwc = layout_name == "NHWC"
    lay = L.NHWC if nhwc else L.NCHW
    xcs, dcs = ((Cin + 3) // 4 * 4, (Cout + 3) // 4 * 4) if nhwc else (0, 0)
    ctx = H.Context(0)
    dev = Dev(ctx)
    try:
        px = dev.put(to_nhwc(x, xcs) if nhwc else x)
        pd = dev.put(to_nhwc(delta, dcs) if nhwc else delta)
        pw = dev.put(w)
        for algo, tol in ((L.ALGO_TC_TF32, 2e-3), (L.ALGO_FP32, 1e-4)):
            pdx = dev.empty(B * Hh * W * (xcs if nhwc else Cin) * 4, fill=0xFF)   # every element must be written
            pdw, pdb = dev.empty(w.nbytes), dev.empty(Cout * 4)
            L.call("dl4jb200_conv_bwd", ctx.handle, C.byref(d), px, lay, xcs, pw, pd, lay, dcs, pdw, pdb, pdx, lay, xcs, algo)
            dx = dev.get(pdx, (B, Hh, W, xcs) if nhwc else x.shape)
            assert np.isfinite(dx).all()
            if nhwc:
                assert not dx[..., Cin:].any()
                dx = from_nhwc(dx, Cin)
            assert rel_err(dx, dxr) <= tol, (algo, rel_err(dx, dxr))
            assert rel_err(dev.get(pdw, w.shape), dwr) <= tol
        # dx alone (what the step driver asks for), result identical to the combined call
        pdx2 = dev.empty(B * Hh * W * (xcs if nhwc else Cin) * 4, fill=0xFF)
        L.call("dl4jb200_conv_bwd", ctx.handle, C.byref(d), px, lay, xcs, pw, pd, lay, dcs, None, None, pdx2, lay, xcs, L.ALGO_FP32)
        np.testing.assert_array_equal(dev.get(pdx2, (B, Hh, W, xcs) if nhwc else x.shape), dev.get(pdx, (B, Hh, W, xcs) if nhwc else x.shape))
    finally:
        dev.free()
        ctx.close()
