"""-m gpu: BatchNormalization (+ReLU) whose apply pass also performs the MAX SubsamplingLayer that consumes it
(dl4jb200_bn_fwd_prestats_maxpool, ResNet-50's stem-batch1 -> stem-maxpool1) against the two-pass path
(dl4jb200_bn_fwd_prestats + dl4jb200_pool_fwd): pooled values AND saved argmax taps bit for bit, same running statistics, with
and without TF32 rounding of stored activations; and against the oracle's BatchNormalization + SubsamplingLayer."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import Dev, from_nhwc, rel_err, to_nhwc

pytestmark = pytest.mark.gpu

CASES = [  # B, C_in, H, W, C (BN channels), pool kernel, pool stride, pool pad, mode, relu
    (4, 64, 30, 30, 64, (3, 3), (2, 2), (0, 0), "Truncate", True),    # the stem's 3x3/2 window (ragged: 30 -> 14)
    (3, 64, 17, 19, 64, (3, 3), (2, 2), (1, 1), "Truncate", True),    # padded border: out-of-range taps never win
    (2, 64, 16, 16, 128, (2, 2), (2, 2), (0, 0), "Truncate", False),  # 2x2/2, identity activation (negative maxima)
    (2, 64, 15, 15, 64, (3, 2), (1, 2), (0, 0), "Same", True),        # generic window (non-square), SAME padding
]


@pytest.mark.parametrize("round_act", [0, 1])
@pytest.mark.parametrize("case", CASES)
def test_bn_maxpool_matches_two_passes(oracle, case, round_act):
    from deeplearning4j_b200 import _lib as L, helpers as H
    B, Cin, Hh, W, Cc, pk, pstr, ppad, mode, relu = case
    rng = np.random.default_rng(5)
    x = rng.standard_normal((B, Cin, Hh, W)).astype(np.float32)
    w = (rng.standard_normal((Cc, Cin, 1, 1)) * 0.3).astype(np.float32)
    b = rng.standard_normal(Cc).astype(np.float32)
    gamma = (1 + 0.1 * rng.standard_normal(Cc)).astype(np.float32)
    beta = (0.1 * rng.standard_normal(Cc)).astype(np.float32)
    d = H.conv_desc(x.shape, Cc, (1, 1), (1, 1), (0, 0), (1, 1), "Truncate")
    pd = H.conv_desc((B, Cc, Hh, W), Cc, pk, pstr, ppad, (1, 1), mode)
    act = L.ACT_RELU if relu else L.ACT_IDENTITY
    ctx = H.Context(0)
    L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_TF32_ACTIVATIONS, round_act)
    dev = Dev(ctx)
    try:
        rows = C.c_int(0)
        L.call("dl4jb200_conv_fwd_bnstats_rows", ctx.handle, C.byref(d), Cin, Cc, L.ALGO_TC_TF32, C.byref(rows))
        assert rows.value > 0
        n, n_out = B * Hh * W * Cc, B * pd.out_h * pd.out_w * Cc
        px, pw, pb, pg, pbt = dev.put(to_nhwc(x)), dev.put(w), dev.put(b), dev.put(gamma), dev.put(beta)
        pz, pst = dev.empty(n * 4), dev.empty(rows.value * 2 * Cc * 4, fill=0xFF)
        piv = dev.put(np.zeros(Cc, np.float32))
        L.call("dl4jb200_conv_fwd_bnstats", ctx.handle, C.byref(d), px, Cin, pw, pb, pz, Cc, L.ALGO_TC_TF32, piv, pst)
        res = []
        for fused in (False, True):
            gm, gv = dev.put(np.zeros(Cc, np.float32)), dev.put(np.ones(Cc, np.float32))
            sm, ss = dev.empty(Cc * 4), dev.empty(Cc * 4)
            pyp, pam = dev.empty(n_out * 4, fill=0xFF), dev.empty(n_out * 4, fill=0xFF)
            if fused:
                L.call("dl4jb200_bn_fwd_prestats_maxpool", ctx.handle, pz, pyp, pam, pg, pbt, gm, gv, sm, ss, B, Cc, Hh, W, Cc, 0.9, 1e-5,
                       1, act, piv, pst, rows.value, C.byref(pd))
            else:
                py = dev.empty(n * 4)
                L.call("dl4jb200_bn_fwd_prestats", ctx.handle, pz, py, pg, pbt, gm, gv, sm, ss, B, Cc, Hh, W, Cc, 0.9, 1e-5, 1, act,
                       None, piv, pst, rows.value)
                L.call("dl4jb200_pool_fwd", ctx.handle, C.byref(pd), py, pyp, pam, L.NHWC, Cc, Cc, L.POOL_MAX, 2)
            res.append([dev.get(pyp, (B, pd.out_h, pd.out_w, Cc)), dev.get(pam, (n_out,), np.int32)] +
                       [dev.get(p_, (Cc,)) for p_ in (gm, gv, sm, ss)])
        for a, f in zip(*res):
            np.testing.assert_array_equal(a, f)
        assert res[1][1].min() >= 0 and res[1][1].max() < pk[0] * pk[1]
        # the oracle's BatchNormalization + activation + SubsamplingLayer on the GPU's own pre-activations
        z = from_nhwc(dev.get(pz, (B, Hh, W, Cc)), Cc)
        yr = oracle.bn_fwd(z, gamma, beta, np.zeros(Cc, np.float32), np.ones(Cc, np.float32), 0.9, 1e-5, True)[0]
        if relu:
            yr = np.maximum(yr, 0)
        pr = oracle.pool_fwd(yr, pk, pstr, ppad, (1, 1), {"Strict": 0, "Truncate": 1, "Same": 2}[mode], 0)
        tol = 2e-3 if round_act else 1e-5
        assert rel_err(from_nhwc(res[1][0], Cc), pr) <= tol, rel_err(from_nhwc(res[1][0], Cc), pr)
    finally:
        dev.free()
        ctx.close()


def test_bn_maxpool_rejects_mismatched_descriptor():
    from deeplearning4j_b200 import _lib as L, helpers as H
    ctx = H.Context(0)
    dev = Dev(ctx)
    try:
        B, Cc, Hh, W = 2, 64, 8, 8
        pd = H.conv_desc((B, Cc, Hh + 1, W), Cc, (2, 2), (2, 2), (0, 0), (1, 1), "Truncate")   # wrong input height
        p = dev.empty(B * Hh * W * Cc * 4)
        v = dev.put(np.ones(Cc, np.float32))
        st = dev.put(np.zeros(2 * Cc, np.float32))
        with pytest.raises(Exception):
            L.call("dl4jb200_bn_fwd_prestats_maxpool", ctx.handle, p, p, None, v, v, v, v, v, v, B, Cc, Hh, W, Cc, 0.9, 1e-5, 1,
                   L.ACT_RELU, v, st, 1, C.byref(pd))
    finally:
        dev.free()
        ctx.close()


BWD_CASES = [  # B, C, H, W, pool kernel, pool pad, relu
    (4, 64, 30, 30, (3, 3), (0, 0), True),     # the stem's window; even size, last row / column outside every window
    (3, 64, 17, 19, (3, 3), (1, 1), True),     # odd sizes (ragged patch grid), padded border
    (2, 128, 16, 16, (2, 2), (0, 0), False),   # non-overlapping windows, identity activation
    (2, 32, 9, 11, (3, 2), (0, 1), True),      # non-square window
]


@pytest.mark.parametrize("round_act", [0, 1])
@pytest.mark.parametrize("case", BWD_CASES)
def test_pool_bwd_bnstats_matches_two_calls(oracle, case, round_act):
    """dl4jb200_pool_bwd_bnstats + dl4jb200_bn_bwd_prestats == dl4jb200_pool_bwd + dl4jb200_bn_bwd: the routed epsilon bit for
    bit, dgamma / dbeta / dx to fp32 summation order; and the oracle's SubsamplingLayer + BatchNormalization backprop."""
    from deeplearning4j_b200 import _lib as L, helpers as H
    import torch
    B, Cc, Hh, W, pk, ppad, relu = case
    rng = np.random.default_rng(11)
    x = (rng.standard_normal((B, Cc, Hh, W)) * 1.3 + 0.2).astype(np.float32)
    gamma = (1 + 0.1 * rng.standard_normal(Cc)).astype(np.float32)
    beta = (0.1 * rng.standard_normal(Cc)).astype(np.float32)
    pd = H.conv_desc((B, Cc, Hh, W), Cc, pk, (2, 2), ppad, (1, 1), "Truncate")
    pe = rng.standard_normal((B, Cc, pd.out_h, pd.out_w)).astype(np.float32)
    act = L.ACT_RELU if relu else L.ACT_IDENTITY
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    ctx = H.Context(0)
    L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_TF32_ACTIVATIONS, round_act)
    dev = Dev(ctx)
    try:
        n, n_out = B * Hh * W * Cc, B * pd.out_h * pd.out_w * Cc
        px, pg, pb, ppe = dev.put(to_nhwc(x)), dev.put(gamma), dev.put(beta), dev.put(to_nhwc(pe))
        gm, gv = dev.put(np.zeros(Cc, np.float32)), dev.put(np.ones(Cc, np.float32))
        sm, ss, py, pyp, pam = dev.empty(Cc * 4), dev.empty(Cc * 4), dev.empty(n * 4), dev.empty(n_out * 4), dev.empty(n_out * 4)
        L.call("dl4jb200_bn_fwd", ctx.handle, px, py, pg, pb, gm, gv, sm, ss, B, Cc, Hh, W, L.NHWC, Cc, 0.9, 1e-5, 1, 1, act, None)
        L.call("dl4jb200_pool_fwd", ctx.handle, C.byref(pd), py, pyp, pam, L.NHWC, Cc, Cc, L.POOL_MAX, 2)
        e0, dx0, dg0, db0 = dev.empty(n * 4, fill=0xFF), dev.empty(n * 4), dev.empty(Cc * 4), dev.empty(Cc * 4)
        L.call("dl4jb200_pool_bwd", ctx.handle, C.byref(pd), py, pyp, ppe, pam, e0, L.NHWC, Cc, Cc, L.POOL_MAX, 2, 1e-8)
        L.call("dl4jb200_bn_bwd", ctx.handle, px, e0, pg, pb, sm, ss, act, dx0, dg0, db0, B, Cc, Hh, W, L.NHWC, Cc)
        e1, dx1, dg1, db1 = dev.empty(n * 4, fill=0xFF), dev.empty(n * 4), dev.empty(Cc * 4), dev.empty(Cc * 4)
        pst, rows = dev.empty(4 * sms * 2 * Cc * 4, fill=0xFF), C.c_int(0)
        L.call("dl4jb200_pool_bwd_bnstats", ctx.handle, C.byref(pd), ppe, pam, e1, Cc, px, pg, pb, sm, ss, int(relu), pst, C.byref(rows))
        assert 0 < rows.value <= 4 * sms
        L.call("dl4jb200_bn_bwd_prestats", ctx.handle, px, e1, pg, pb, sm, ss, act, dx1, dg1, db1, B, Cc, Hh, W, Cc, pst, rows.value)
        np.testing.assert_array_equal(dev.get(e1, (n,)), dev.get(e0, (n,)))
        a0, a1 = dev.get(dx0, (B, Hh, W, Cc)), dev.get(dx1, (B, Hh, W, Cc))
        assert rel_err(a1, a0) <= (1e-3 if round_act else 2e-6), rel_err(a1, a0)
        assert rel_err(dev.get(dg1, (Cc,)), dev.get(dg0, (Cc,))) <= 1e-5
        assert rel_err(dev.get(db1, (Cc,)), dev.get(db0, (Cc,))) <= 1e-5
        y = from_nhwc(dev.get(py, (B, Hh, W, Cc)), Cc)
        e_full = oracle.pool_bwd(y, pe, pk, (2, 2), ppad, (1, 1), oracle.TRUNCATE, oracle.POOL_MAX)
        if relu:
            e_full = np.where(y > 0, e_full, 0).astype(np.float32)
        dxr, dgr, dbr = oracle.bn_bwd(x, e_full, gamma, dev.get(sm, (Cc,)), dev.get(ss, (Cc,)))
        assert rel_err(from_nhwc(a1, Cc), dxr) <= (2e-3 if round_act else 2e-5)
        tol_sum = 1e-3 if round_act else 1e-4
        assert rel_err(dev.get(dg1, (Cc,)), dgr) <= tol_sum and rel_err(dev.get(db1, (Cc,)), dbr) <= tol_sum
    finally:
        dev.free()
        ctx.close()


def test_pool_bwd_bnstats_declines_other_windows():
    from deeplearning4j_b200 import _lib as L, helpers as H
    ctx = H.Context(0)
    dev = Dev(ctx)
    try:
        B, Cc, Hh, W = 2, 64, 9, 9
        pd = H.conv_desc((B, Cc, Hh, W), Cc, (3, 3), (1, 1), (0, 0), (1, 1), "Truncate")   # stride 1
        p = dev.empty(B * Hh * W * Cc * 4, fill=0)
        v = dev.put(np.ones(Cc, np.float32))
        rows = C.c_int(0)
        with pytest.raises(Exception, match="stride"):
            L.call("dl4jb200_pool_bwd_bnstats", ctx.handle, C.byref(pd), p, p, p, Cc, p, v, v, v, v, 1, p, C.byref(rows))
    finally:
        dev.free()
        ctx.close()
