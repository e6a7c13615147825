"""-m gpu: backward batch-norm sums taken in the epilogue of the backprop-data GEMM that produces the layer's epsilon
(dl4jb200_conv_bwd_data_bnstats + dl4jb200_bn_bwd_prestats) against the separate-pass path (conv_bwd / conv_bwd_data_fused
+ bn_bwd) and the oracle: the epsilon written by the GEMM is bit-identical, dgamma / dbeta / dx agree to fp32 summation order."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import Dev, from_nhwc, rel_err, to_nhwc

pytestmark = pytest.mark.gpu

CASES = [  # B, C (BN channels = conv c_in), H, W, conv c_out, k, mode, bn_relu, fused (prior + relu mask of the owning Add)
    (8, 64, 28, 28, 64, 3, "Same", True, False),      # bn2a <- backprop-data of the 3x3
    (8, 128, 14, 14, 512, 1, "Truncate", True, False),  # bn2b <- backprop-data of the 1x1 expansion
    (8, 256, 14, 14, 64, 1, "Truncate", False, True),   # bn2c through its Add: fan-in sum + ReLU backprop in the same epilogue
    (4, 1024, 7, 7, 256, 1, "Truncate", False, True),
]


@pytest.mark.parametrize("case", CASES)
def test_bwd_bnstats_matches_separate_pass(oracle, case):
    from deeplearning4j_b200 import _lib as L, helpers as H
    o = oracle
    B, Cc, Hh, W, Cout, k, mode, bn_relu, fused = case
    rng = np.random.default_rng(17)
    x = (rng.standard_normal((B, Cc, Hh, W)) * 1.5 + 0.3).astype(np.float32)     # the BatchNormalization's input
    w = (rng.standard_normal((Cout, Cc, k, k)) * 0.1).astype(np.float32)
    gamma, beta = (rng.random(Cc) + 0.5).astype(np.float32), (rng.standard_normal(Cc) * 0.3).astype(np.float32)
    d = H.conv_desc((B, Cc, Hh, W), Cout, (k, k), (1, 1), (0, 0), (1, 1), mode)
    delta = rng.standard_normal((B, Cout, d.out_h, d.out_w)).astype(np.float32)
    prior = rng.standard_normal(x.shape).astype(np.float32) if fused else None
    act_out = rng.standard_normal(x.shape).astype(np.float32) if fused else None     # the Add vertex' (post-ReLU) output
    ctx = H.Context(0)
    L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_TF32_ACTIVATIONS, 1)
    dev = Dev(ctx)
    try:
        n = B * Hh * W * Cc
        px, pw, pd = dev.put(to_nhwc(x)), dev.put(w), dev.put(to_nhwc(delta))
        pg, pb = dev.put(gamma), dev.put(beta)
        gm, gv = dev.put(np.zeros(Cc, np.float32)), dev.put(np.ones(Cc, np.float32))
        sm, ss, py = dev.empty(Cc * 4), dev.empty(Cc * 4), dev.empty(n * 4)
        act = L.ACT_RELU if bn_relu else L.ACT_IDENTITY
        L.call("dl4jb200_bn_fwd", ctx.handle, px, py, pg, pb, gm, gv, sm, ss, B, Cc, Hh, W, L.NHWC, Cc, 0.9, 1e-5, 1, 1, act, None)
        pp = dev.put(to_nhwc(prior)) if fused else None
        pa = dev.put(to_nhwc(act_out)) if fused else None
        # reference path: (fused) backprop-data, then bn_bwd with its own partial-sum pass
        pe0, pdx0, pdg0, pdb0 = dev.empty(n * 4), dev.empty(n * 4), dev.empty(Cc * 4), dev.empty(Cc * 4)
        L.call("dl4jb200_conv_bwd_data_fused", ctx.handle, C.byref(d), pw, pd, Cout, pe0, Cc, pp, pa, L.ACT_RELU if fused else L.ACT_IDENTITY,
               L.ALGO_TC_TF32)
        L.call("dl4jb200_bn_bwd", ctx.handle, px, pe0, pg, pb, sm, ss, act, pdx0, pdg0, pdb0, B, Cc, Hh, W, L.NHWC, Cc)
        # sums in the epilogue
        rows = C.c_int(0)
        L.call("dl4jb200_conv_bwd_data_bnstats_rows", ctx.handle, C.byref(d), Cout, Cc, L.ALGO_TC_TF32, C.byref(rows))
        assert rows.value > 0
        pst = dev.empty(rows.value * 2 * Cc * 4, fill=0xFF)
        pe1, pdx1, pdg1, pdb1 = dev.empty(n * 4), dev.empty(n * 4), dev.empty(Cc * 4), dev.empty(Cc * 4)
        L.call("dl4jb200_conv_bwd_data_bnstats", ctx.handle, C.byref(d), pw, pd, Cout, pe1, Cc, pp, pa, L.ACT_RELU if fused else L.ACT_IDENTITY,
               L.ALGO_TC_TF32, px, pg, pb, sm, ss, int(bn_relu), pst)
        L.call("dl4jb200_bn_bwd_prestats", ctx.handle, px, pe1, pg, pb, sm, ss, act, pdx1, pdg1, pdb1, B, Cc, Hh, W, Cc, pst, rows.value)
        e0, e1 = dev.get(pe0, (B, Hh, W, Cc)), dev.get(pe1, (B, Hh, W, Cc))
        np.testing.assert_array_equal(e1, e0)                       # the epsilon itself is untouched by the extra epilogue work
        assert np.isfinite(dev.get(pst, (rows.value * 2 * Cc,))).all()  # every partial row was written
        for a_, b_ in ((pdg1, pdg0), (pdb1, pdb0)):
            assert rel_err(dev.get(a_, (Cc,)), dev.get(b_, (Cc,))) <= 2e-5
        # dx is stored rounded to TF32 (DL4JB200_OPT_TF32_ACTIVATIONS is on, as in the training step): coefficients that differ
        # in the last fp32 bit may round a value to the neighbouring TF32 number, 2^-10 of that value
        assert rel_err(dev.get(pdx1, (B, Hh, W, Cc)), dev.get(pdx0, (B, Hh, W, Cc))) <= 1e-3
        # and against the oracle, fed the epsilon the GEMM produced
        mean, std = dev.get(sm, (Cc,)), dev.get(ss, (Cc,))
        e = from_nhwc(e1, Cc)
        if bn_relu:
            y = (x - mean[None, :, None, None]) * (gamma / std)[None, :, None, None] + beta[None, :, None, None]
            e = np.where(y > 0, e, 0).astype(np.float32)
        dxr, dgr, dbr = o.bn_bwd(x, e, gamma, mean, std)
        assert rel_err(dev.get(pdg1, (Cc,)), dgr) <= 1e-4 and rel_err(dev.get(pdb1, (Cc,)), dbr) <= 1e-4
        assert rel_err(from_nhwc(dev.get(pdx1, (B, Hh, W, Cc)), Cc), dxr) <= 6e-4  # one rounding to nearest TF32 (2^-11)
    finally:
        dev.free()
        ctx.close()
