"""-m gpu: the layers that share ConvolutionLayer's helper field -- Deconvolution2D and
SeparableConvolution2D -- through the C ABI against the oracle restatement of their libnd4j ops."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import Dev, from_nhwc, rel_err, to_nhwc

pytestmark = pytest.mark.gpu

DECONV_CASES = [  # B, nIn, H, W, nOut, k, s, p, d
    (2, 8, 5, 6, 12, (3, 3), (2, 2), (1, 1), (1, 1)),
    (3, 32, 7, 7, 16, (2, 2), (2, 2), (0, 0), (1, 1)),      # the classic x2 up-sampling deconvolution
    (2, 5, 6, 4, 7, (3, 2), (1, 2), (0, 1), (2, 1)),        # ragged channels, dilation, asymmetric geometry
    (4, 64, 8, 8, 32, (4, 4), (2, 2), (1, 1), (1, 1)),      # TMA-eligible channel counts
]


@pytest.mark.parametrize("case", DECONV_CASES)
@pytest.mark.parametrize("algo_name", ["fp32", "tf32"])
@pytest.mark.parametrize("layout_name", ["NCHW", "NHWC"])
def test_deconvolution2d_matches_oracle(oracle, case, algo_name, layout_name):
    from deeplearning4j_b200 import _lib as L, helpers as H
    o = oracle
    B, nIn, Hh, W, nOut, k, s, p, d = case
    rng = np.random.default_rng(11)
    x = rng.standard_normal((B, nIn, Hh, W)).astype(np.float32)
    w = (rng.standard_normal((nIn, nOut) + k) * 0.3).astype(np.float32)
    b = rng.standard_normal(nOut).astype(np.float32)
    Ho, Wo = o.deconv_out_size(Hh, k[0], s[0], p[0], d[0]), o.deconv_out_size(W, k[1], s[1], p[1], d[1])
    delta = rng.standard_normal((B, nOut, Ho, Wo)).astype(np.float32)
    algo, tol = (L.ALGO_FP32, 1e-4) if algo_name == "fp32" else (L.ALGO_TC_TF32, 2e-3)
    # (the TF32 bound is relative to max|z| of the pre-activation, as for the plain convolution; a squashing activation
    # would rescale it, so the fused tanh is exercised in the FP32 run)
    act = L.ACT_TANH if algo_name == "fp32" else L.ACT_IDENTITY
    nhwc = layout_name == "NHWC"
    lay = L.NHWC if nhwc else L.NCHW
    xcs, zcs = ((nIn + 3) // 4 * 4, (nOut + 3) // 4 * 4) if nhwc else (0, 0)
    desc = L.ConvDesc(B, nIn, Hh, W, nOut, k[0], k[1], s[0], s[1], p[0], p[1], d[0], d[1], Ho, Wo)
    ctx = H.Context(0)
    dev = Dev(ctx)
    try:
        px = dev.put(to_nhwc(x, xcs) if nhwc else x)
        pd = dev.put(to_nhwc(delta, zcs) if nhwc else delta)
        pw, pb = dev.put(w), dev.put(b)
        nz = B * Ho * Wo * (zcs if nhwc else nOut)
        nx = B * Hh * W * (xcs if nhwc else nIn)
        pz, pdx, pdw, pdb = dev.empty(nz * 4), dev.empty(nx * 4), dev.empty(w.nbytes), dev.empty(nOut * 4)
        L.call("dl4jb200_deconv2d_fwd", ctx.handle, C.byref(desc), px, lay, xcs, pw, pb, pz, lay, zcs, act, algo)
        L.call("dl4jb200_deconv2d_bwd", ctx.handle, C.byref(desc), px, lay, xcs, pw, pd, lay, zcs, pdw, pdb, pdx, lay, xcs, algo)
        z = dev.get(pz, (B, Ho, Wo, zcs) if nhwc else (B, nOut, Ho, Wo))
        dx = dev.get(pdx, (B, Hh, W, xcs) if nhwc else x.shape)
        if nhwc:
            assert not z[..., nOut:].any() and not dx[..., nIn:].any()  # padding channels stay zero
            z, dx = from_nhwc(z, nOut), from_nhwc(dx, nIn)
        zr = o.deconv2d_fwd(x, w, b, s, p, d).astype(np.float64)
        zr = np.tanh(zr) if act == L.ACT_TANH else zr
        dwr, dbr, dxr = o.deconv2d_bwd(x, w, delta, s, p, d)
        assert rel_err(z, zr) <= tol
        assert rel_err(dev.get(pdw, w.shape), dwr) <= tol
        assert rel_err(dev.get(pdb, (nOut,)), dbr) <= 1e-5
        assert rel_err(dx, dxr) <= tol
    finally:
        dev.free()
        ctx.close()


SEP_CASES = [  # B, nIn, H, W, mult, nOut, k, s, pad, d, mode
    (2, 8, 9, 9, 1, 16, (3, 3), (1, 1), (0, 0), (1, 1), "Same"),
    (3, 6, 10, 8, 2, 10, (3, 3), (2, 2), (1, 1), (1, 1), "Truncate"),
    (2, 32, 7, 7, 1, 64, (5, 5), (1, 1), (2, 2), (1, 1), "Truncate"),
    (2, 5, 8, 8, 3, 7, (3, 2), (1, 2), (0, 0), (2, 1), "Same"),
]


@pytest.mark.parametrize("case", SEP_CASES)
def test_separable_convolution2d_matches_oracle(oracle, case):
    from deeplearning4j_b200 import _lib as L, helpers as H
    o = oracle
    B, nIn, Hh, W, mult, nOut, k, s, pad, d, mode = case
    rng = np.random.default_rng(13)
    x = rng.standard_normal((B, nIn, Hh, W)).astype(np.float32)
    wd = (rng.standard_normal((mult, nIn) + k) * 0.4).astype(np.float32)
    wp = (rng.standard_normal((nOut, nIn * mult, 1, 1)) * 0.3).astype(np.float32)
    b = rng.standard_normal(nOut).astype(np.float32)
    omode = {"Same": o.SAME, "Truncate": o.TRUNCATE}[mode]
    Ho, Wo, pT, pL = o.conv_geometry(Hh, W, k, s, pad, d, omode)
    zr, ur = o.sepconv2d_fwd(x, wd, wp, b, s, pad, d, omode)
    delta = rng.standard_normal(zr.shape).astype(np.float32)
    dwdr, dwpr, dbr, dxr = o.sepconv2d_bwd(x, wd, wp, delta, s, pad, d, omode)
    Cm = nIn * mult
    xcs, ucs, zcs = (nIn + 3) // 4 * 4, (Cm + 3) // 4 * 4, (nOut + 3) // 4 * 4
    dd = L.ConvDesc(B, nIn, Hh, W, Cm, k[0], k[1], s[0], s[1], pT, pL, d[0], d[1], Ho, Wo)
    dp = L.ConvDesc(B, Cm, Ho, Wo, nOut, 1, 1, 1, 1, 0, 0, 1, 1, Ho, Wo)
    ctx = H.Context(0)
    dev = Dev(ctx)
    try:
        px, pwd, pwp, pb = dev.put(to_nhwc(x, xcs)), dev.put(wd), dev.put(wp), dev.put(b)
        pdl = dev.put(to_nhwc(delta, zcs))
        pu, pz = dev.empty(B * Ho * Wo * ucs * 4), dev.empty(B * Ho * Wo * zcs * 4)
        pdu, pdx = dev.empty(B * Ho * Wo * ucs * 4), dev.empty(B * Hh * W * xcs * 4)
        pdwd, pdwp, pdb = dev.empty(wd.nbytes), dev.empty(wp.nbytes), dev.empty(nOut * 4)
        # preOutput: depth-wise, then the point-wise 1x1 convolution (+ bias) on the tensor cores' FP32 twin
        L.call("dl4jb200_depthwise_fwd", ctx.handle, C.byref(dd), mult, px, xcs, pwd, pu, ucs)
        L.call("dl4jb200_conv_fwd", ctx.handle, C.byref(dp), pu, L.NHWC, ucs, pwp, pb, pz, L.NHWC, zcs, L.ACT_IDENTITY, L.ALGO_FP32)
        # backprop: point-wise conv_bwd gives dWp, db and the epsilon of u; depth-wise backward gives dWd and dx
        L.call("dl4jb200_conv_bwd", ctx.handle, C.byref(dp), pu, L.NHWC, ucs, pwp, pdl, L.NHWC, zcs, pdwp, pdb, pdu, L.NHWC, ucs, L.ALGO_FP32)
        L.call("dl4jb200_depthwise_bwd", ctx.handle, C.byref(dd), mult, px, xcs, pwd, pdu, ucs, pdwd, pdx, xcs)
        u = dev.get(pu, (B, Ho, Wo, ucs))
        assert not u[..., Cm:].any()
        assert rel_err(from_nhwc(u, Cm), ur) <= 1e-5
        assert rel_err(from_nhwc(dev.get(pz, (B, Ho, Wo, zcs)), nOut), zr) <= 1e-4
        assert rel_err(dev.get(pdwp, wp.shape), dwpr) <= 1e-4
        assert rel_err(dev.get(pdb, (nOut,)), dbr) <= 1e-5
        assert rel_err(dev.get(pdwd, wd.shape), dwdr) <= 1e-4
        dx = dev.get(pdx, (B, Hh, W, xcs))
        assert not dx[..., nIn:].any()
        assert rel_err(from_nhwc(dx, nIn), dxr) <= 1e-4
    finally:
        dev.free()
        ctx.close()
