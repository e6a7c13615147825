"""-m gpu parity tests: the CUDA path (through the C ABI, via the helper-SPI mirror) against the CPU oracle on
identical seeded inputs -- the pattern of deeplearning4j-cuda/src/test/.../TestConvolution.java:71-151
(helper vs built-in) -- plus the reference's golden vectors replayed through the GPU path.

Tolerances: pooling / BN / LRN fp32 kernels rel <= 1e-5 of max|ref|; pooling argmax and max
values bit-exact; convolution FP32 mode <= 1e-4 (the reference's own helper-vs-builtin bound,
TestConvolution.java:145,151), TF32 tensor-core mode <= 2e-3 of max|ref|.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import load_golden
from gpu_util import Dev, from_nhwc, rel_err, to_nhwc

pytestmark = pytest.mark.gpu

TOL_BW = 1e-5
TOL_CONV_FP32 = 1e-4
TOL_CONV_TF32 = 2e-3
TOL_CONV_TF32_QUANTISED = 1e-5  # TF32 path vs the oracle on identically quantised operands (fp32 accumulation order only)


@pytest.fixture(scope="module")
def H():
    from deeplearning4j_b200 import build
    build.build()
    from deeplearning4j_b200 import helpers
    return helpers


@pytest.fixture(scope="module")
def ctx(H):
    c = H.Context(0)
    yield c
    c.close()


def _a(v, shape):
    return np.asarray(v, np.float32).reshape(shape)


MODES = {0: "Strict", 1: "Truncate", 2: "Same"}

CONV_CASES = [  # B, Cin, H, W, Cout, k, s, p, d, mode
    (2, 3, 7, 8, 4, (3, 3), (1, 1), (1, 1), (1, 1), 1),
    (2, 3, 8, 9, 5, (3, 3), (2, 2), (0, 0), (1, 1), 2),
    (1, 2, 9, 9, 3, (3, 3), (1, 1), (0, 0), (2, 2), 2),
    (2, 4, 6, 6, 3, (1, 1), (2, 2), (0, 0), (1, 1), 1),
    (2, 1, 12, 12, 2, (5, 5), (2, 2), (2, 2), (1, 1), 1),
    (3, 1, 28, 28, 20, (5, 5), (1, 1), (0, 0), (1, 1), 2),      # LeNet conv1 (SAME)
    (3, 20, 14, 14, 50, (5, 5), (1, 1), (0, 0), (1, 1), 2),     # LeNet conv2: Cout=50 (not a multiple of 4)
    (2, 3, 38, 38, 64, (7, 7), (2, 2), (0, 0), (1, 1), 1),      # ResNet-50 stem shape class (K=7*7*4)
    (2, 64, 9, 9, 64, (1, 1), (2, 2), (0, 0), (1, 1), 1),       # ResNet-50 strided 1x1
    (2, 64, 7, 7, 256, (1, 1), (1, 1), (0, 0), (1, 1), 1),      # 1x1 expand, N=256 -> two 128-column tiles
    (2, 128, 7, 7, 128, (3, 3), (1, 1), (0, 0), (1, 1), 2),     # 3x3 SAME, K=1152
    (5, 3, 23, 23, 64, (11, 11), (4, 4), (2, 2), (1, 1), 1),    # AlexNet conv1 class
    (2, 192, 7, 7, 136, (3, 3), (1, 1), (0, 0), (1, 1), 2),     # ragged N (136 = 128 + 8)
    (1, 8, 4, 4, 8, (3, 3), (1, 1), (0, 0), (1, 1), 2),         # M=16 << one tile
    # Convolution1DLayer reshapes [B,C,T] to [B,C,T,1] and runs the inherited 2-D code with kernel (k,1), stride (s,1),
    # pad (p,0) (nn/layers/convolution/Convolution1DLayer.java): the same helper serves it
    (3, 16, 50, 1, 24, (3, 1), (1, 1), (1, 0), (1, 1), 1),
    (2, 32, 41, 1, 32, (5, 1), (2, 1), (0, 0), (1, 1), 2),
    (2, 64, 33, 1, 64, (3, 1), (1, 1), (0, 0), (2, 1), 2),      # dilated, SAME, TMA-eligible channel count
    # DenseLayer / OutputLayer shapes run as 1x1 convolutions on a 1x1 map: M = minibatch (one row tile), long K -> the split-K
    # form of the GEMM (partial accumulators per K range, ordered fold with bias), ragged N = 1000; backprop-data N = 2048
    (64, 2048, 1, 1, 1000, (1, 1), (1, 1), (0, 0), (1, 1), 1),
    (130, 1024, 1, 1, 1024, (1, 1), (1, 1), (0, 0), (1, 1), 1),  # two row tiles (one ragged), split-K in both directions
    # 1x1 windows whose channel counts are not multiples of 32 still take the TMA path (single tap: the ragged last K-chunk
    # is zero-filled): forward C = 40, backprop-data C = 72; and LeNet's dense shapes 800 -> 500 -> 10
    (3, 40, 9, 9, 72, (1, 1), (1, 1), (0, 0), (1, 1), 1),
    (64, 800, 1, 1, 500, (1, 1), (1, 1), (0, 0), (1, 1), 1),
    (64, 500, 1, 1, 10, (1, 1), (1, 1), (0, 0), (1, 1), 1),
]


@pytest.mark.parametrize("algo_name", ["fp32", "tf32"])
@pytest.mark.parametrize("case", CONV_CASES, ids=[str(i) for i in range(len(CONV_CASES))])
def test_conv_forward_backward_vs_oracle(H, ctx, oracle, case, algo_name):
    B, Cin, Hh, W, Cout, k, s, p, d, mode = case
    algo = H.ALGO_FP32 if algo_name == "fp32" else H.ALGO_TC_TF32
    tol = TOL_CONV_FP32 if algo_name == "fp32" else TOL_CONV_TF32
    rng = np.random.default_rng(12345)
    x = rng.random((B, Cin, Hh, W), dtype=np.float32)  # U[0,1) like BenchmarkDataSetIterator
    w = (rng.standard_normal((Cout, Cin) + k) * 0.5).astype(np.float32)
    b = rng.standard_normal(Cout).astype(np.float32)
    helper = H.CudnnConvolutionHelper(ctx, algo=algo)
    assert helper.checkSupported()
    z = helper.preOutput(x, w, b, k, s, p, None, None, MODES[mode], d)
    zr = oracle.conv_fwd(x, w, b, stride=s, pad=p, dilation=d, mode=mode)
    assert z.shape == zr.shape
    assert rel_err(z, zr) <= tol, ("z", rel_err(z, zr))
    delta = rng.standard_normal(zr.shape).astype(np.float32)
    flat = np.full(Cout + w.size, np.nan, np.float32)  # bias first, then W: ConvolutionParamInitializer.java:116-123
    bview, wview = flat[:Cout].reshape(1, Cout), flat[Cout:].reshape(w.shape)
    grad, eps_next = helper.backpropGradient(x, w, delta, k, s, p, bview, wview, "identity", None, None, None,
                                             MODES[mode], d)
    dwr, dbr, dxr = oracle.conv_bwd(x, w, delta, stride=s, pad=p, dilation=d, mode=mode)
    assert grad["W"] is wview and not np.isnan(flat).any()  # written in place into the flat gradient vector
    assert rel_err(wview, dwr) <= tol, ("dW", rel_err(wview, dwr))
    assert rel_err(bview.ravel(), dbr) <= max(tol, 1e-5), ("db", rel_err(bview.ravel(), dbr))
    assert rel_err(eps_next, dxr) <= tol, ("dx", rel_err(eps_next, dxr))
    if algo_name == "tf32":
        # The whole TF32 deviation above is OPERAND QUANTISATION: filters are rounded to nearest TF32 when packed, the
        # tensor core ignores the low 13 mantissa bits of activations / deltas. Against the oracle fed operands
        # quantised the same way the tcgen05 path is an fp32-accumulate convolution to ~1e-6.
        from oracle.net_oracle import tf32_round, tf32_trunc
        zq = oracle.conv_fwd(tf32_trunc(x), tf32_round(w), b, stride=s, pad=p, dilation=d, mode=mode)
        assert rel_err(z, zq) <= TOL_CONV_TF32_QUANTISED, ("z vs quantised oracle", rel_err(z, zq))
        dwq, _, dxq = oracle.conv_bwd(tf32_trunc(x), tf32_round(w), tf32_trunc(delta), stride=s, pad=p, dilation=d, mode=mode)
        assert rel_err(wview, dwq) <= TOL_CONV_TF32_QUANTISED, ("dW vs quantised oracle", rel_err(wview, dwq))
        assert rel_err(eps_next, dxq) <= TOL_CONV_TF32_QUANTISED, ("dx vs quantised oracle", rel_err(eps_next, dxq))


def test_conv_golden_contained(H, ctx):
    # ConvolutionLayerTest.java:264-278 replayed through the helper: preOutput then activate(sigmoid)
    g = load_golden("conv_contained")
    x, w, b = _a(g["input"], g["input_shape"]), _a(g["w"], g["w_shape"]), _a(g["b"], [1, 2])
    for algo in (H.ALGO_FP32, H.ALGO_TC_TF32):  # 0.5 and small integers are exact in TF32
        helper = H.CudnnConvolutionHelper(ctx, algo=algo)
        z = helper.preOutput(x, w, b, g["kernel"], g["stride"], g["pad"], None, None, "Truncate", (1, 1))
        a = helper.activate(z, "sigmoid")
        assert a is z
        np.testing.assert_allclose(a, _a(g["out"], g["out_shape"]), rtol=0, atol=1e-6)
    assert helper.activate(z, "softsign") is None  # unsupported activation => null => layer falls back
    assert helper.preOutput(x.astype(np.float64), w.astype(np.float64), None, g["kernel"], g["stride"], g["pad"]) is None


def test_conv_no_bias_and_fused_activation(H, ctx, oracle):
    from deeplearning4j_b200 import _lib
    rng = np.random.default_rng(5)
    x = rng.standard_normal((2, 8, 6, 6)).astype(np.float32)
    w = rng.standard_normal((16, 8, 3, 3)).astype(np.float32) * 0.2
    zr = oracle.conv_fwd(x, w, None, stride=(1, 1), pad=(1, 1))
    d = H.conv_desc(x.shape, 16, (3, 3), (1, 1), (1, 1), (1, 1), "Truncate")
    dev = Dev(ctx)
    try:
        for algo, tol in ((H.ALGO_FP32, TOL_CONV_FP32), (H.ALGO_TC_TF32, TOL_CONV_TF32)):
            for act, kind in (("relu", oracle.ACT_RELU), ("tanh", oracle.ACT_TANH), ("sigmoid", oracle.ACT_SIGMOID)):
                px, pw = dev.put(to_nhwc(x)), dev.put(w)
                pz = dev.empty(2 * 6 * 6 * 16 * 4)
                _lib.call("dl4jb200_conv_fwd", ctx.handle, C.byref(d), px, H.NHWC, 8, pw, None, pz, H.NHWC, 16,
                          H.ACTIVATIONS[act], algo)
                got = from_nhwc(dev.get(pz, (2, 6, 6, 16)), 16)
                assert rel_err(got, oracle.act_fwd(zr, kind)) <= tol
    finally:
        dev.free()


def test_tf32_activation_rounding_option(H, ctx, oracle):
    """DL4JB200_OPT_TF32_ACTIVATIONS: producers store round-to-nearest TF32 values (== tf32_round of the plain output);
    a fused-activation convolution does too, a plain preOutput (feeds BatchNormalization) does not."""
    from deeplearning4j_b200 import _lib
    from oracle.net_oracle import tf32_round
    rng = np.random.default_rng(11)
    x = rng.standard_normal((2, 8, 6, 6)).astype(np.float32)
    w = (rng.standard_normal((16, 8, 3, 3)) * 0.2).astype(np.float32)
    d = H.conv_desc(x.shape, 16, (3, 3), (1, 1), (1, 1), (1, 1), "Truncate")
    dev = Dev(ctx)
    n = x.size

    def run(opt):
        _lib.call("dl4jb200_ctx_set_option", ctx.handle, _lib.OPT_TF32_ACTIVATIONS, opt)
        out = {}
        px, pw = dev.put(x), dev.put(w)
        pn = dev.empty(n * 4)
        _lib.call("dl4jb200_nchw_to_nhwc", ctx.handle, px, pn, 2, 8, 6, 6, 8)
        out["relayout"] = dev.get(pn, (n,))
        pa = dev.empty(n * 4)
        _lib.call("dl4jb200_act_fwd", ctx.handle, px, pa, n, H.ACTIVATIONS["tanh"])
        out["tanh"] = dev.get(pa, (n,))
        ps = dev.empty(n * 4)
        _lib.call("dl4jb200_add_act_fwd", ctx.handle, px, pw, ps, n, H.ACTIVATIONS["relu"])
        out["add_relu"] = dev.get(ps, (n,))
        for act in ("identity", "relu"):
            pz = dev.empty(2 * 6 * 6 * 16 * 4)
            _lib.call("dl4jb200_conv_fwd", ctx.handle, C.byref(d), pn, H.NHWC, 8, pw, None, pz, H.NHWC, 16,
                      H.ACTIVATIONS[act], H.ALGO_TC_TF32)
            out["conv_" + act] = dev.get(pz, (2 * 6 * 6 * 16,))
        return out

    try:
        plain, rounded = run(0), run(1)
        for k in ("relayout", "tanh", "add_relu"):
            np.testing.assert_array_equal(rounded[k], tf32_round(plain[k]))
            assert (plain[k].view(np.uint32) & 0x1FFF).any()
        assert not (rounded["conv_relu"].view(np.uint32) & 0x1FFF).any()
        assert (rounded["conv_identity"].view(np.uint32) & 0x1FFF).any()
        with pytest.raises(Exception):
            _lib.call("dl4jb200_ctx_set_option", ctx.handle, 99, 1)
    finally:
        _lib.call("dl4jb200_ctx_set_option", ctx.handle, _lib.OPT_TF32_ACTIVATIONS, 0)
        dev.free()


@pytest.mark.parametrize("case", [
    # B, Cin, H, W, Cout, k, s, p, mode
    (4, 64, 14, 14, 64, (3, 3), (1, 1), (1, 1), 2),      # TMA persistent kernel, fused epilogue, full tiles
    (3, 256, 7, 7, 64, (1, 1), (1, 1), (0, 0), 1),   # 1x1, ragged last M tile (147 rows)
    (2, 40, 9, 9, 24, (3, 3), (1, 1), (1, 1), 2),        # channel counts that are not tile multiples
    (2, 64, 14, 14, 128, (1, 1), (2, 2), (0, 0), 1),  # strided 1x1: compact walk, elementwise fallback
    (64, 2048, 1, 1, 1024, (1, 1), (1, 1), (0, 0), 1),  # dense-shaped: split-K GEMM, the fused epilogue runs in its fold
])
@pytest.mark.parametrize("algo_name", ["fp32", "tf32"])
def test_conv_bwd_data_fused_epilogue(H, ctx, oracle, case, algo_name):
    """dx = relu'(act_out) * (conv_bwd_data(delta) + eps_sum): separate prior buffer, in-place prior, no prior, no mask."""
    from deeplearning4j_b200 import _lib
    B, Cin, Hh, W, Cout, k, s, p, mode = case
    algo = H.ALGO_FP32 if algo_name == "fp32" else H.ALGO_TC_TF32
    tol = TOL_CONV_FP32 if algo_name == "fp32" else TOL_CONV_TF32
    rng = np.random.default_rng(77)
    x = rng.standard_normal((B, Cin, Hh, W)).astype(np.float32)
    w = (rng.standard_normal((Cout, Cin) + k) * 0.2).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, k, s, p, (1, 1), MODES[mode])
    delta = rng.standard_normal((B, Cout, d.out_h, d.out_w)).astype(np.float32)
    prior = rng.standard_normal(x.shape).astype(np.float32)
    act_out = np.maximum(rng.standard_normal(x.shape), 0).astype(np.float32)  # a ReLU output: ~half zeros
    _, _, dxr = oracle.conv_bwd(x, w, delta, stride=s, pad=p, dilation=(1, 1), mode=mode)
    cs_in, cs_out = (Cin + 3) // 4 * 4, (Cout + 3) // 4 * 4
    dev = Dev(ctx)
    try:
        pw, pd = dev.put(w), dev.put(to_nhwc(delta))
        pp, pm = dev.put(to_nhwc(prior)), dev.put(to_nhwc(act_out))
        n = B * Hh * W * cs_in

        def run(eps_sum, mask, inplace=False):
            pdx = dev.put(to_nhwc(prior)) if inplace else dev.empty(n * 4, fill=0)
            _lib.call("dl4jb200_conv_bwd_data_fused", ctx.handle, C.byref(d), pw, pd, cs_out, pdx, cs_in,
                      pdx if inplace else eps_sum, pm if mask else None, H.ACTIVATIONS["relu" if mask else "identity"], algo)
            return from_nhwc(dev.get(pdx, (B, Hh, W, cs_in)), Cin)

        scale = max(np.abs(dxr).max(), np.abs(prior).max())
        for eps_sum, mask, inplace in ((pp, True, False), (None, True, True), (None, True, False), (pp, False, False), (None, False, False)):
            want = dxr + (prior if (eps_sum is not None or inplace) else 0)
            if mask:
                want = want * (act_out > 0)
            got = run(eps_sum, mask, inplace)
            assert np.abs(got - want).max() <= tol * scale, (eps_sum is not None, mask, inplace, np.abs(got - want).max() / scale)
            if mask:
                assert not got[act_out <= 0].any()
        # the elementwise form of the same bookkeeping
        pa, pb_ = dev.put(to_nhwc(dxr.astype(np.float32))), dev.put(to_nhwc(prior))
        po = dev.empty(n * 4)
        _lib.call("dl4jb200_eps_sum_act_bwd", ctx.handle, pa, pb_, pm, po, n, H.ACTIVATIONS["relu"])
        got = from_nhwc(dev.get(po, (B, Hh, W, cs_in)), Cin)
        np.testing.assert_array_equal(got, ((dxr.astype(np.float32) + prior) * (act_out > 0)).astype(np.float32))
    finally:
        dev.free()


@pytest.mark.parametrize("case", [(4, 64, 14, 14, 64, (3, 3), 2), (3, 256, 7, 7, 96, (1, 1), 1), (2, 32, 9, 9, 200, (3, 3), 2)])
def test_conv_bias_gradient_rides_in_filter_gemm(H, ctx, oracle, case):
    """With DL4JB200_OPT_TF32_ACTIVATIONS the TF32 filter-gradient GEMM also produces db (delta^T x ones as an extra
    accumulator column); it must equal the plain column sum of delta, and dW must be unchanged."""
    from deeplearning4j_b200 import _lib
    from oracle.net_oracle import tf32_round
    B, Cin, Hh, W, Cout, k, mode = case
    rng = np.random.default_rng(3)
    x = tf32_round(rng.random((B, Cin, Hh, W), dtype=np.float32))
    w = (rng.standard_normal((Cout, Cin) + k) * 0.3).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, k, (1, 1), (k[0] // 2, k[1] // 2), (1, 1), MODES[mode])
    delta = tf32_round(rng.standard_normal((B, Cout, d.out_h, d.out_w)).astype(np.float32))
    dwr, dbr, _ = oracle.conv_bwd(x, w, delta, stride=(1, 1), pad=(k[0] // 2, k[1] // 2), dilation=(1, 1), mode=mode)
    cs_in, cs_out = (Cin + 3) // 4 * 4, (Cout + 3) // 4 * 4
    dev = Dev(ctx)
    try:
        px, pw, pd = dev.put(to_nhwc(x)), dev.put(w), dev.put(to_nhwc(delta))
        res = {}
        for opt in (0, 1):
            _lib.call("dl4jb200_ctx_set_option", ctx.handle, _lib.OPT_TF32_ACTIVATIONS, opt)
            pdw, pdb = dev.empty(w.size * 4, fill=0), dev.empty(Cout * 4, fill=0)
            _lib.call("dl4jb200_conv_bwd", ctx.handle, C.byref(d), px, H.NHWC, cs_in, pw, pd, H.NHWC, cs_out, pdw, pdb, None,
                      H.NHWC, cs_in, H.ALGO_TC_TF32)
            res[opt] = (dev.get(pdw, w.shape), dev.get(pdb, (Cout,)))
        np.testing.assert_array_equal(res[0][0], res[1][0])
        for opt in (0, 1):
            assert rel_err(res[opt][1], dbr) <= 1e-5, (opt, rel_err(res[opt][1], dbr))
            assert rel_err(res[opt][0], dwr) <= TOL_CONV_TF32_QUANTISED
    finally:
        _lib.call("dl4jb200_ctx_set_option", ctx.handle, _lib.OPT_TF32_ACTIVATIONS, 0)
        dev.free()


def test_prepacked_filters_match_per_call_packing(H, oracle):
    """filter_register + filters_repack: registered filters give bit-identical conv_fwd / conv_bwd / fused results, follow
    the parameters after a repack, and fall back to per-call packing after filters_invalidate."""
    from deeplearning4j_b200 import _lib
    ctx = H.Context(0)  # own context: the registry lives in it
    rng = np.random.default_rng(21)
    B, Cin, Hh, W, Cout, k = 2, 64, 10, 10, 96, (3, 3)
    x = rng.random((B, Cin, Hh, W), dtype=np.float32)
    w = (rng.standard_normal((Cout, Cin) + k) * 0.2).astype(np.float32)
    w2 = (rng.standard_normal((Cout, Cin) + k) * 0.2).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, k, (1, 1), (1, 1), (1, 1), "Same")
    delta = rng.standard_normal((B, Cout, d.out_h, d.out_w)).astype(np.float32)
    dev = Dev(ctx)
    try:
        px, pw, pd = dev.put(to_nhwc(x)), dev.put(w), dev.put(to_nhwc(delta))
        launches = _lib.cdll().dl4jb200_launch_count

        def run():
            pz, pdx = dev.empty(B * Hh * W * Cout * 4), dev.empty(B * Hh * W * Cin * 4)
            n0 = launches()
            _lib.call("dl4jb200_conv_fwd", ctx.handle, C.byref(d), px, H.NHWC, Cin, pw, None, pz, H.NHWC, Cout, 0, H.ALGO_TC_TF32)
            _lib.call("dl4jb200_conv_bwd", ctx.handle, C.byref(d), px, H.NHWC, Cin, pw, pd, H.NHWC, Cout, None, None, pdx, H.NHWC, Cin,
                      H.ALGO_TC_TF32)
            return dev.get(pz, (B, Hh, W, Cout)), dev.get(pdx, (B, Hh, W, Cin)), launches() - n0

        z0, dx0, l0 = run()
        _lib.call("dl4jb200_filter_register", ctx.handle, C.byref(d), pw, Cin, Cout, H.ALGO_TC_TF32, 1)
        z1, dx1, l1 = run()                      # registered but never packed: still per-call
        assert l1 == l0
        _lib.call("dl4jb200_filters_repack", ctx.handle)
        z2, dx2, l2 = run()
        assert l2 == l0 - 2                      # both pack kernels gone
        for a, b in ((z0, z1), (z0, z2), (dx0, dx1), (dx0, dx2)):
            np.testing.assert_array_equal(a, b)
        _lib.call("dl4jb200_memcpy_h2d", ctx.handle, pw, w2.ctypes.data_as(C.c_void_p), w2.nbytes)  # "the updater ran"
        _lib.call("dl4jb200_filters_repack", ctx.handle)
        z3, dx3, _ = run()
        zr = oracle.conv_fwd(x, w2, None, stride=(1, 1), pad=(1, 1), dilation=(1, 1), mode=2)
        assert rel_err(from_nhwc(z3, Cout), zr) <= TOL_CONV_TF32
        assert np.abs(z3 - z0).max() > 0.1
        _lib.call("dl4jb200_filters_invalidate", ctx.handle)
        z4, dx4, l4 = run()
        assert l4 == l0
        np.testing.assert_array_equal(z3, z4)
        np.testing.assert_array_equal(dx3, dx4)
        with pytest.raises(Exception):
            _lib.call("dl4jb200_filter_register", ctx.handle, C.byref(d), pw, Cin, Cout, H.ALGO_TC_TF32, 1)
    finally:
        dev.free()
        ctx.close()


@pytest.mark.parametrize("case", [
    (8, 64, 14, 14, 64, (3, 3), 2),      # one 64-wide N tile, 13 M tiles
    (4, 64, 14, 14, 256, (1, 1), 1),     # two 128-wide N tiles
    (3, 256, 7, 7, 512, (1, 1), 1),      # four N tiles, ragged last M tile (147 rows)
    (64, 128, 28, 28, 128, (3, 3), 2),   # 392 M tiles: every CTA accumulates over several tiles before the flush
])
def test_conv_epilogue_bn_statistics(H, ctx, oracle, case):
    """conv_fwd_bnstats + bn_fwd_prestats == conv_fwd + bn_fwd (and the oracle): same z bit for bit, same normalised output,
    saved statistics and running statistics, for a pivot far from the batch mean as well as for a good one."""
    from deeplearning4j_b200 import _lib
    B, Cin, Hh, W, Cout, k, mode = case
    rng = np.random.default_rng(17)
    x = rng.random((B, Cin, Hh, W), dtype=np.float32)
    w = (rng.standard_normal((Cout, Cin) + k) * 0.3).astype(np.float32)
    b = rng.standard_normal(Cout).astype(np.float32)
    gamma = (1 + 0.1 * rng.standard_normal(Cout)).astype(np.float32)
    beta = (0.1 * rng.standard_normal(Cout)).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, k, (1, 1), (k[0] // 2, k[1] // 2), (1, 1), MODES[mode])
    rows = C.c_int(0)
    _lib.call("dl4jb200_conv_fwd_bnstats_rows", ctx.handle, C.byref(d), Cin, Cout, H.ALGO_TC_TF32, C.byref(rows))
    assert rows.value > 0
    n_out = B * d.out_h * d.out_w * Cout
    dev = Dev(ctx)
    try:
        px, pw, pb, pg, pbt = dev.put(to_nhwc(x)), dev.put(w), dev.put(b), dev.put(gamma), dev.put(beta)
        # reference path: conv_fwd then bn_fwd
        pz0, py0 = dev.empty(n_out * 4), dev.empty(n_out * 4)
        gm0, gv0 = dev.put(np.zeros(Cout, np.float32)), dev.put(np.ones(Cout, np.float32))
        sm0, ss0 = dev.empty(Cout * 4), dev.empty(Cout * 4)
        _lib.call("dl4jb200_conv_fwd", ctx.handle, C.byref(d), px, H.NHWC, Cin, pw, pb, pz0, H.NHWC, Cout, 0, H.ALGO_TC_TF32)
        _lib.call("dl4jb200_bn_fwd", ctx.handle, pz0, py0, pg, pbt, gm0, gv0, sm0, ss0, B, Cout, d.out_h, d.out_w, H.NHWC, Cout,
                  0.9, 1e-5, 1, 1, H.ACTIVATIONS["relu"], None)
        z0, y0 = dev.get(pz0, (n_out,)), dev.get(py0, (n_out,))
        ref = [dev.get(p_, (Cout,)) for p_ in (gm0, gv0, sm0, ss0)]
        zmean = z0.reshape(-1, Cout).mean(0)
        for pivot_kind in ("zero", "near"):
            pivot = np.zeros(Cout, np.float32) if pivot_kind == "zero" else (zmean + 0.05).astype(np.float32)
            pz, py, pst = dev.empty(n_out * 4), dev.empty(n_out * 4), dev.empty(rows.value * 2 * Cout * 4, fill=0xFF)
            # the pivot buffer doubles as the running mean here only for the "zero" case (gmean starts at 0)
            gm = dev.put(np.zeros(Cout, np.float32))
            pp = gm if pivot_kind == "zero" else dev.put(pivot)
            gv, sm, ss = dev.put(np.ones(Cout, np.float32)), dev.empty(Cout * 4), dev.empty(Cout * 4)
            _lib.call("dl4jb200_conv_fwd_bnstats", ctx.handle, C.byref(d), px, Cin, pw, pb, pz, Cout, H.ALGO_TC_TF32, pp, pst)
            _lib.call("dl4jb200_bn_fwd_prestats", ctx.handle, pz, py, pg, pbt, gm, gv, sm, ss, B, Cout, d.out_h, d.out_w, Cout,
                      0.9, 1e-5, 1, H.ACTIVATIONS["relu"], None, pp, pst, rows.value)
            np.testing.assert_array_equal(dev.get(pz, (n_out,)), z0)
            assert rel_err(dev.get(py, (n_out,)), y0) <= 2e-5, pivot_kind
            for got_p, want in zip((gm, gv, sm, ss), ref):
                assert rel_err(dev.get(got_p, (Cout,)), want) <= 1e-5, pivot_kind
    finally:
        dev.free()
    # a convolution that cannot carry them says so
    d2 = H.conv_desc((2, 3, 9, 9), 5, (3, 3), (1, 1), (1, 1), (1, 1), "Truncate")
    _lib.call("dl4jb200_conv_fwd_bnstats_rows", ctx.handle, C.byref(d2), 4, 8, H.ALGO_TC_TF32, C.byref(rows))
    assert rows.value == 0


def test_conv_invalid_shapes_raise_like_reference(H, ctx):
    from deeplearning4j_b200 import DL4JInvalidConfigException, DL4JInvalidInputException
    helper = H.CudnnConvolutionHelper(ctx)
    x = np.zeros((1, 1, 10, 10), np.float32)
    w = np.zeros((1, 1, 3, 3), np.float32)
    with pytest.raises(DL4JInvalidConfigException):  # TestConvolutionModes.java:33-110: Strict, 10x10, k3 s3
        helper.preOutput(x, w, None, (3, 3), (3, 3), (0, 0), None, None, "Strict", (1, 1))
    with pytest.raises(DL4JInvalidInputException):
        helper.preOutput(np.zeros((1, 1, 2, 2), np.float32), w, None, (3, 3), (1, 1), (0, 0), None, None, "Truncate", (1, 1))


# ------------------------------------------------------------------ pooling
def test_subsampling_golden(H, ctx):
    # SubsamplingLayerTest.java:47-133 through the helper; the max-backprop case contains ties
    g = load_golden("subsampling")
    x = _a(g["input"], g["input_shape"])
    h = H.CudnnSubsamplingHelper(ctx)
    y = h.activate(x, True, g["kernel"], g["stride"], g["pad"], "MAX", "Truncate", (1, 1))
    np.testing.assert_array_equal(y, _a(g["max_out"], g["out_shape"]))
    np.testing.assert_array_equal(h.argmax()[0, 1], [[2, 1], [0, 0]])
    _, dx = h.backpropGradient(x, _a(g["max_bwd_eps"], g["out_shape"]), g["kernel"], g["stride"], g["pad"], "MAX")
    np.testing.assert_array_equal(dx, _a(g["max_bwd_out"], g["input_shape"]))
    y = h.activate(x, True, g["kernel"], g["stride"], g["pad"], "AVG")
    np.testing.assert_array_equal(y, _a(g["avg_out"], g["out_shape"]))
    _, dx = h.backpropGradient(x, _a(g["avg_bwd_eps"], g["out_shape"]), g["kernel"], g["stride"], g["pad"], "AVG")
    np.testing.assert_array_equal(dx, _a(g["avg_bwd_out"], g["input_shape"]))
    assert h.backpropGradient(x, _a(g["avg_bwd_eps"], g["out_shape"]), g["kernel"], g["stride"], g["pad"], "SUM") is None


POOL_CASES = [  # B, C, H, W, k, s, p, d, mode
    (2, 3, 9, 10, (3, 3), (2, 2), (1, 1), (1, 1), 1),
    (2, 5, 12, 12, (3, 3), (2, 2), (0, 0), (1, 1), 1),   # ResNet-50 stem pool class
    (2, 4, 7, 7, (3, 3), (2, 2), (0, 0), (1, 1), 2),     # SAME, asymmetric padding
    (1, 2, 14, 14, (3, 3), (7, 7), (0, 0), (1, 1), 2),   # AlexNet pool3 class (stride 7)
    (2, 3, 8, 8, (2, 2), (2, 2), (0, 0), (1, 1), 0),
    (1, 2, 11, 11, (3, 3), (1, 1), (0, 0), (2, 2), 1),   # dilated
    (2, 8, 4, 4, (3, 3), (2, 2), (0, 0), (1, 1), 1),     # ResNet-50 head pool 4x4 -> 1x1
    (2, 6, 31, 1, (3, 1), (2, 1), (0, 0), (1, 1), 1),    # Subsampling1DLayer: [B,C,T] as [B,C,T,1], kernel (k,1)
    (2, 8, 20, 1, (2, 1), (2, 1), (0, 0), (1, 1), 2),    # stride (2,1): width stride 1 keeps it off the 2x2-patch kernel
    (3, 8, 16, 16, (2, 2), (2, 2), (0, 0), (1, 1), 1),   # 2x2/2 (VGG / LeNet class) on the 2x2-patch backward kernel
]


@pytest.mark.parametrize("layout", ["NCHW", "NHWC"])
@pytest.mark.parametrize("case", POOL_CASES, ids=[str(i) for i in range(len(POOL_CASES))])
def test_pool_vs_oracle(H, ctx, oracle, case, layout):
    from deeplearning4j_b200 import _lib
    B, Cc, Hh, W, k, s, p, d, mode = case
    rng = np.random.default_rng(99)
    # quantised values => plenty of ties inside windows: exercises the first-max rule
    x = (rng.integers(-3, 4, (B, Cc, Hh, W))).astype(np.float32) * 0.5
    desc = H.conv_desc(x.shape, Cc, k, s, p, d, MODES[mode])
    yshape = (B, Cc, desc.out_h, desc.out_w)
    eps = rng.standard_normal(yshape).astype(np.float32)
    nh = layout == "NHWC"
    cs = (Cc + 3) // 4 * 4 if nh else 0
    lay = H.NHWC if nh else H.NCHW
    enc = (lambda a: to_nhwc(a)) if nh else (lambda a: a)
    dec = (lambda a, shp: from_nhwc(a.reshape(shp[0], shp[2], shp[3], cs), shp[1])) if nh else (lambda a, shp: a.reshape(shp))
    n_y = B * desc.out_h * desc.out_w * (cs if nh else Cc)
    n_x = B * Hh * W * (cs if nh else Cc)
    dev = Dev(ctx)
    try:
        px, pe = dev.put(enc(x)), dev.put(enc(eps))
        for ptype, oname in ((H.POOL_MAX, oracle.POOL_MAX), (H.POOL_AVG, oracle.POOL_AVG), (H.POOL_PNORM, oracle.POOL_PNORM)):
            py, pa, pdx = dev.empty(n_y * 4), dev.empty(n_y * 4), dev.empty(n_x * 4)
            _lib.call("dl4jb200_pool_fwd", ctx.handle, C.byref(desc), px, py, pa, lay, cs, cs, ptype, 2)
            y = dec(dev.get(py, (n_y,)), yshape)
            if ptype == H.POOL_MAX:
                yr, amr = oracle.pool_fwd(x, k, s, p, d, mode, oname, want_argmax=True)
                np.testing.assert_array_equal(y, yr)  # bit-exact
                am = dev.get(pa, (n_y,), np.int32)
                am = from_nhwc(am.reshape(B, desc.out_h, desc.out_w, cs), Cc) if nh else am.reshape(yshape)
                np.testing.assert_array_equal(am, amr)  # bit-exact indices, first max wins
            else:
                yr = oracle.pool_fwd(x, k, s, p, d, mode, oname, pnorm=2)
                assert rel_err(y, yr) <= TOL_BW
            dxr = oracle.pool_bwd(x, eps, k, s, p, d, mode, oname, pnorm=2)
            for use_saved in ((True, False) if ptype == H.POOL_MAX else (False,)):
                _lib.call("dl4jb200_pool_bwd", ctx.handle, C.byref(desc), px, py, pe, pa if use_saved else None, pdx,
                          lay, cs, cs, ptype, 2, 1e-8)
                dx = dec(dev.get(pdx, (n_x,)), x.shape)
                assert rel_err(dx, dxr) <= TOL_BW, (ptype, use_saved, rel_err(dx, dxr))
    finally:
        dev.free()


# ------------------------------------------------------------------ batch norm
@pytest.mark.parametrize("shape", [(4, 3, 5, 6), (8, 64, 14, 14), (3, 20, 7, 9), (2, 260, 4, 4), (16, 8, 1, 1)])
def test_batchnorm_helper_vs_oracle(H, ctx, oracle, shape):
    rng = np.random.default_rng(12345)
    B, Cc, Hh, W = shape
    x = (rng.standard_normal(shape) * 3 + 2).astype(np.float32)
    gamma = (rng.random(Cc) + 0.5).astype(np.float32)
    beta = rng.standard_normal(Cc).astype(np.float32)
    # params view layout gamma|beta|mean|var (BatchNormalizationParamInitializer.java:80-113)
    flat = np.concatenate([gamma, beta, np.zeros(Cc, np.float32), np.ones(Cc, np.float32)])
    gm, gv = flat[2 * Cc:3 * Cc].reshape(1, Cc), flat[3 * Cc:].reshape(1, Cc)
    gm_r, gv_r = np.zeros(Cc, np.float32), np.ones(Cc, np.float32)
    h = H.CudnnBatchNormalizationHelper(ctx)
    assert h.checkSupported(1e-5)
    for it in range(2):  # two iterations: running statistics accumulate
        y = h.preOutput(x, True, list(shape), gamma.reshape(1, Cc), beta.reshape(1, Cc), gm, gv, 0.9, 1e-5)
        yr, m_r, s_r = oracle.bn_fwd(x, gamma, beta, gm_r, gv_r, 0.9, 1e-5, True)
        assert rel_err(y, yr) <= TOL_BW
        np.testing.assert_allclose(gm.ravel(), gm_r, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(gv.ravel(), gv_r, rtol=1e-5, atol=1e-6)  # var + eps is what gets stored
    e = rng.standard_normal(shape).astype(np.float32)
    gflat = np.full(4 * Cc, np.nan, np.float32)
    dg, db = gflat[:Cc].reshape(1, Cc), gflat[Cc:2 * Cc].reshape(1, Cc)
    grad, dx = h.backpropGradient(x, e, list(shape), gamma.reshape(1, Cc), dg, db, 1e-5)
    dxr, dgr, dbr = oracle.bn_bwd(x, e, gamma, m_r, s_r)
    assert rel_err(dx, dxr) <= 2e-5, rel_err(dx, dxr)
    assert rel_err(dg.ravel(), dgr) <= TOL_BW and rel_err(db.ravel(), dbr) <= TOL_BW
    # inference: sqrt(gVar) with no eps added (BatchNormalization.java:274-279); running stats untouched
    before = flat.copy()
    yt = h.preOutput(x, False, list(shape), gamma.reshape(1, Cc), beta.reshape(1, Cc), gm, gv, 0.9, 1e-5)
    ytr, _, _ = oracle.bn_fwd(x, gamma, beta, gm_r.copy(), gv_r.copy(), 0.9, 1e-5, False)
    assert rel_err(yt, ytr) <= TOL_BW
    np.testing.assert_array_equal(flat, before)


def test_batchnorm_mean0_std1_property(H, ctx):
    # BatchNormalizationTest.java:178-208
    rng = np.random.default_rng(1)
    x = (rng.standard_normal((10, 6, 9, 9)) * 5 + 11).astype(np.float32)
    h = H.CudnnBatchNormalizationHelper(ctx)
    gm, gv = np.zeros((1, 6), np.float32), np.ones((1, 6), np.float32)
    y = h.preOutput(x, True, [10, 6, 9, 9], np.ones((1, 6), np.float32), np.zeros((1, 6), np.float32), gm, gv, 0.9, 1e-5)
    np.testing.assert_allclose(y.astype(np.float64).mean((0, 2, 3)), 0, atol=1e-5)
    np.testing.assert_allclose(y.astype(np.float64).std((0, 2, 3)), 1, atol=1e-4)


@pytest.mark.parametrize("shape", [(4, 8, 5, 6), (8, 64, 14, 14), (3, 50, 7, 7), (2, 2048, 4, 4)])
def test_batchnorm_nhwc_fused_relu(H, ctx, oracle, shape):
    from deeplearning4j_b200 import _lib
    rng = np.random.default_rng(3)
    B, Cc, Hh, W = shape
    cs = (Cc + 3) // 4 * 4
    x = (rng.standard_normal(shape) * 2 + 1).astype(np.float32)
    gamma = (rng.random(Cc) + 0.5).astype(np.float32)
    beta = rng.standard_normal(Cc).astype(np.float32)
    e = rng.standard_normal(shape).astype(np.float32)
    gm_r, gv_r = np.zeros(Cc, np.float32), np.ones(Cc, np.float32)
    yr, m_r, s_r = oracle.bn_fwd(x, gamma, beta, gm_r, gv_r, 0.9, 1e-5, True)
    ar = oracle.act_fwd(yr, oracle.ACT_RELU)
    dxr, dgr, dbr = oracle.bn_bwd(x, oracle.act_bwd(yr, e, oracle.ACT_RELU), gamma, m_r, s_r)
    dev = Dev(ctx)
    try:
        px, pe = dev.put(to_nhwc(x)), dev.put(to_nhwc(e))
        pg, pb = dev.put(gamma), dev.put(beta)
        pm, pv = dev.put(np.zeros(Cc, np.float32)), dev.put(np.ones(Cc, np.float32))
        n = B * Hh * W * cs
        py, pdx = dev.empty(n * 4), dev.empty(n * 4)
        psm, pss, pdg, pdb = (dev.empty(Cc * 4) for _ in range(4))
        _lib.call("dl4jb200_bn_fwd", ctx.handle, px, py, pg, pb, pm, pv, psm, pss, B, Cc, Hh, W, H.NHWC, cs, 0.9, 1e-5, 1, 1,
                  H.ACT_RELU, None)
        a = dev.get(py, (B, Hh, W, cs))
        assert rel_err(from_nhwc(a, Cc), ar) <= TOL_BW
        assert not a[..., Cc:].any()  # padding channels stay zero
        np.testing.assert_allclose(dev.get(pm, (Cc,)), gm_r, rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(dev.get(pv, (Cc,)), gv_r, rtol=1e-5, atol=1e-6)
        _lib.call("dl4jb200_bn_bwd", ctx.handle, px, pe, pg, pb, psm, pss, H.ACT_RELU, pdx, pdg, pdb, B, Cc, Hh, W, H.NHWC, cs)
        # elements with BN output exactly 0 are measure-zero for random data; masks agree
        assert rel_err(from_nhwc(dev.get(pdx, (B, Hh, W, cs)), Cc), dxr) <= 2e-5
        assert rel_err(dev.get(pdg, (Cc,)), dgr) <= TOL_BW and rel_err(dev.get(pdb, (Cc,)), dbr) <= TOL_BW
    finally:
        dev.free()


# ------------------------------------------------------------------ LRN
def test_lrn_golden_and_oracle(H, ctx, oracle):
    g = load_golden("lrn")
    x, e = _a(g["x"], g["shape"]), _a(g["eps"], g["shape"])
    h = H.CudnnLocalResponseNormalizationHelper(ctx)
    assert h.checkSupported(g["k"], g["n"], g["alpha"], g["beta"])
    y = h.activate(x, True, g["k"], g["n"], g["alpha"], g["beta"])
    np.testing.assert_allclose(y, _a(g["y"], g["shape"]), rtol=0, atol=1e-5)  # LocalResponseTest.java:36-60
    _, dx = h.backpropGradient(x, e, g["k"], g["n"], g["alpha"], g["beta"])
    exp = _a(g["dx"], g["shape"]).ravel()
    for i in g["dx_asserted_indices"]:  # :110-117
        assert abs(dx.ravel()[i] - exp[i]) <= g["dx_tol"]
    assert rel_err(dx, oracle.lrn_bwd(x, e, g["k"], g["n"], g["alpha"], g["beta"])) <= TOL_BW


@pytest.mark.parametrize("layout", ["NCHW", "NHWC"])
def test_lrn_vs_oracle(H, ctx, oracle, layout):
    from deeplearning4j_b200 import _lib
    rng = np.random.default_rng(8)
    B, Cc, Hh, W = 3, 18, 5, 7
    x = rng.standard_normal((B, Cc, Hh, W)).astype(np.float32) * 3
    e = rng.standard_normal(x.shape).astype(np.float32)
    nh = layout == "NHWC"
    cs = 20 if nh else 0
    dev = Dev(ctx)
    try:
        for (k, n, alpha, beta) in ((2.0, 5.0, 1e-4, 0.75), (1.0, 4.0, 1e-2, 0.5), (2.0, 1.0, 1e-3, 0.75)):
            px = dev.put(to_nhwc(x, cs) if nh else x)
            pe = dev.put(to_nhwc(e, cs) if nh else e)
            n_el = B * Hh * W * (cs if nh else Cc)
            py, pdx = dev.empty(n_el * 4), dev.empty(n_el * 4)
            lay = H.NHWC if nh else H.NCHW
            _lib.call("dl4jb200_lrn_fwd", ctx.handle, px, py, B, Cc, Hh, W, lay, cs, k, n, alpha, beta)
            _lib.call("dl4jb200_lrn_bwd", ctx.handle, px, pe, pdx, B, Cc, Hh, W, lay, cs, k, n, alpha, beta)
            y, dx = dev.get(py, (n_el,)), dev.get(pdx, (n_el,))
            if nh:
                y, dx = from_nhwc(y.reshape(B, Hh, W, cs), Cc), from_nhwc(dx.reshape(B, Hh, W, cs), Cc)
            else:
                y, dx = y.reshape(x.shape), dx.reshape(x.shape)
            assert rel_err(y, oracle.lrn_fwd(x, k, n, alpha, beta)) <= TOL_BW
            assert rel_err(dx, oracle.lrn_bwd(x, e, k, n, alpha, beta)) <= TOL_BW
    finally:
        dev.free()


# ------------------------------------------------------------------ callers either side (§8f)
def test_elementwise_softmax_updater(H, ctx, oracle):
    from deeplearning4j_b200 import _lib
    rng = np.random.default_rng(21)
    dev = Dev(ctx)
    try:
        n = 4099  # odd: exercises the scalar path
        for nn in (n, 4096):
            a, b = rng.standard_normal(nn).astype(np.float32), rng.standard_normal(nn).astype(np.float32)
            pa, pb, po = dev.put(a), dev.put(b), dev.empty(nn * 4)
            _lib.call("dl4jb200_add_act_fwd", ctx.handle, pa, pb, po, nn, H.ACT_RELU)
            np.testing.assert_array_equal(dev.get(po, (nn,)), np.maximum(a + b, 0))
            for name, kind in (("relu", oracle.ACT_RELU), ("sigmoid", oracle.ACT_SIGMOID), ("tanh", oracle.ACT_TANH),
                               ("leakyrelu", oracle.ACT_LEAKYRELU), ("identity", oracle.ACT_IDENTITY)):
                _lib.call("dl4jb200_act_fwd", ctx.handle, pa, po, nn, H.ACTIVATIONS[name])
                assert rel_err(dev.get(po, (nn,)), oracle.act_fwd(a, kind)) <= TOL_BW
                _lib.call("dl4jb200_act_bwd", ctx.handle, pa, pb, po, nn, H.ACTIVATIONS[name])
                assert rel_err(dev.get(po, (nn,)), oracle.act_bwd(a, b, kind)) <= TOL_BW
            _lib.call("dl4jb200_accumulate", ctx.handle, pa, pb, nn)
            np.testing.assert_array_equal(dev.get(pa, (nn,)), a + b)
        # softmax + NLL, labels = column 1 (BenchmarkDataSetIterator.java:132-139)
        B, K = 37, 1000
        z = rng.standard_normal((B, K)).astype(np.float32) * 4
        lab = np.zeros((B, K), np.float32)
        lab[:, 1] = 1
        pz, pl, pd, ploss = dev.put(z), dev.put(lab), dev.empty(B * K * 4), dev.empty(4)
        _lib.call("dl4jb200_softmax_nll", ctx.handle, pz, K, pl, pd, K, ploss, B, K)
        loss_r, _, delta_r = oracle.softmax_nll(z, lab)
        assert rel_err(dev.get(pd, (B, K)), delta_r) <= TOL_BW
        assert abs(dev.get(ploss, (1,))[0] - loss_r) <= 1e-5 * abs(loss_r)
        # updaters (TestUpdaters.java formulas), two steps each, with l1/l2 and minibatch division
        m = 1000
        g = rng.standard_normal(m).astype(np.float32)
        w0 = rng.standard_normal(m).astype(np.float32)
        for kind, okind, ns, args in ((H._lib.UPD_RMSPROP, oracle.UPD_RMSPROP, 1, dict(lr=0.1, h1=0.96, h2=1e-3)),
                                      (H._lib.UPD_NESTEROVS, oracle.UPD_NESTEROVS, 1, dict(lr=1e-2, h1=0.9)),
                                      (H._lib.UPD_ADADELTA, oracle.UPD_ADADELTA, 2, dict(h1=0.95, h2=1e-6)),
                                      (H._lib.UPD_SGD, oracle.UPD_SGD, 1, dict(lr=0.05)),
                                      (H._lib.UPD_NOOP, oracle.UPD_NOOP, 1, dict())):
            wr, sr = w0.copy(), np.zeros(ns * m, np.float32)
            pw, pg, ps = dev.put(w0), dev.put(g), dev.put(np.zeros(ns * m, np.float32))
            for _ in range(2):
                oracle.updater_apply(wr, g, sr, okind, l1=1e-7, l2=5e-5, batch_div=16.0, **args)
                _lib.call("dl4jb200_updater_apply", ctx.handle, pw, pg, ps, m, kind, args.get("lr", 0.0), args.get("h1", 0.0),
                          args.get("h2", 0.0), 1e-7, 5e-5, 16.0)
            assert rel_err(dev.get(pw, (m,)), wr) <= TOL_BW, kind
    finally:
        dev.free()


def test_relayout_roundtrip(H, ctx):
    from deeplearning4j_b200 import _lib
    rng = np.random.default_rng(4)
    dev = Dev(ctx)
    try:
        for shape, cs in (((2, 3, 5, 7), 4), ((3, 50, 7, 7), 52), ((2, 64, 9, 9), 64), ((1, 33, 33, 31), 36)):
            x = rng.standard_normal(shape).astype(np.float32)
            B, Cc, Hh, W = shape
            px, pn, pb = dev.put(x), dev.empty(B * Hh * W * cs * 4, fill=0xFF), dev.empty(x.nbytes)
            _lib.call("dl4jb200_nchw_to_nhwc", ctx.handle, px, pn, B, Cc, Hh, W, cs)
            np.testing.assert_array_equal(dev.get(pn, (B, Hh, W, cs)), to_nhwc(x, cs))
            _lib.call("dl4jb200_nhwc_to_nchw", ctx.handle, pn, pb, B, Cc, Hh, W, cs)
            np.testing.assert_array_equal(dev.get(pb, shape), x)
    finally:
        dev.free()


@pytest.mark.parametrize("case", [(4, 3, 32, 32, 64, 7, 2, 3), (2, 3, 30, 26, 16, 7, 2, 3), (3, 1, 28, 28, 20, 5, 2, 2), (2, 4, 17, 19, 40, 3, 2, 1),
                                  (2, 3, 24, 28, 64, 3, 1, 1),    # VGG16-style 3x3 stride 1
                                  (3, 1, 28, 28, 20, 5, 1, 2),    # LeNet's 5x5 stride 1
                                  (2, 3, 33, 29, 32, 7, 4, 2),    # stride 4
                                  (2, 2, 19, 23, 24, 5, 3, 0),    # stride 3, no padding
                                  (2, 3, 59, 63, 64, 11, 4, 2),   # AlexNet's 11x11 stride 4: two 8-pixel K-chunks per filter row
                                  (2, 3, 40, 44, 32, 9, 2, 4),    # 9 taps, stride 2 (second chunk two windows... four starts right)
                                  (3, 4, 21, 25, 16, 15, 1, 7)])  # the widest row: 15 taps + the dummy fill both chunks
def test_rowpacked_stem_conv_vs_oracle(H, ctx, oracle, case):
    """dl4jb200_conv_rowpack_fwd / _wgrad (TMA-fed stem path) against the oracle's ordinary convolution."""
    from deeplearning4j_b200 import _lib
    B, Cin, Hh, W, Cout, k, s, p = case
    rng = np.random.default_rng(77)
    x = rng.random((B, Cin, Hh, W), dtype=np.float32)
    w = (rng.standard_normal((Cout, Cin, k, k)) * 0.3).astype(np.float32)
    b = rng.standard_normal(Cout).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, (k, k), (s, s), (p, p), (1, 1), "Truncate")
    zr = oracle.conv_fwd(x, w, b, stride=(s, s), pad=(p, p))
    delta = rng.standard_normal(zr.shape).astype(np.float32)
    dwr, dbr, _ = oracle.conv_bwd(x, w, delta, stride=(s, s), pad=(p, p), need_dx=False)
    pitch = max(s * (d.out_w - 1) + 8 * ((k + 1 + 7) // 8), p + 1 + W)
    pitch += pitch % 2
    zcs = (Cout + 3) // 4 * 4
    dev = Dev(ctx)
    try:
        px = dev.put(x)
        pxp = dev.empty(B * Hh * pitch * 4 * 4, fill=0)
        _lib.call("dl4jb200_nchw_to_rowpack4", ctx.handle, px, pxp, B, Cin, Hh, W, pitch, p + 1)
        pw, pb = dev.put(w), dev.put(b)
        pz = dev.empty(B * d.out_h * d.out_w * zcs * 4)
        _lib.call("dl4jb200_conv_rowpack_fwd", ctx.handle, C.byref(d), pxp, pitch, pw, pb, pz, zcs, H.ACT_IDENTITY)
        z = from_nhwc(dev.get(pz, (B, d.out_h, d.out_w, zcs)), Cout)
        assert rel_err(z, zr) <= TOL_CONV_TF32, rel_err(z, zr)
        pd = dev.put(to_nhwc(delta, zcs))
        pdw, pdb = dev.empty(w.nbytes), dev.empty(Cout * 4)
        _lib.call("dl4jb200_conv_rowpack_wgrad", ctx.handle, C.byref(d), pxp, pitch, pd, zcs, pdw, None)
        _lib.call("dl4jb200_bias_grad", ctx.handle, pd, B * d.out_h * d.out_w, zcs, Cout, pdb)
        assert rel_err(dev.get(pdw, w.shape), dwr) <= TOL_CONV_TF32
        assert rel_err(dev.get(pdb, (Cout,)), dbr) <= 1e-5
        if Cout % 32 == 0:  # the stem forward can also carry the following BatchNormalization's statistics
            import torch
            sms = torch.cuda.get_device_properties(0).multi_processor_count
            gamma, beta = np.ones(Cout, np.float32), np.zeros(Cout, np.float32)
            outs = []
            for fusedstats in (False, True):
                pz2, py = dev.empty(B * d.out_h * d.out_w * zcs * 4), dev.empty(B * d.out_h * d.out_w * zcs * 4)
                gm, gv = dev.put(np.zeros(Cout, np.float32)), dev.put(np.ones(Cout, np.float32))
                sm_, ss_ = dev.empty(Cout * 4), dev.empty(Cout * 4)
                pg_, pbt_ = dev.put(gamma), dev.put(beta)
                if fusedstats:
                    pst, rows = dev.empty(4 * sms * 2 * zcs * 4), C.c_int(0)
                    _lib.call("dl4jb200_conv_rowpack_fwd_bnstats", ctx.handle, C.byref(d), pxp, pitch, pw, pb, pz2, zcs, gm, pst, C.byref(rows))
                    assert rows.value > 0
                    _lib.call("dl4jb200_bn_fwd_prestats", ctx.handle, pz2, py, pg_, pbt_, gm, gv, sm_, ss_, B, Cout, d.out_h, d.out_w, zcs,
                              0.9, 1e-5, 1, H.ACT_IDENTITY, None, gm, pst, rows.value)
                else:
                    _lib.call("dl4jb200_conv_rowpack_fwd", ctx.handle, C.byref(d), pxp, pitch, pw, pb, pz2, zcs, H.ACT_IDENTITY)
                    _lib.call("dl4jb200_bn_fwd", ctx.handle, pz2, py, pg_, pbt_, gm, gv, sm_, ss_, B, Cout, d.out_h, d.out_w, H.NHWC, zcs,
                              0.9, 1e-5, 1, 1, H.ACT_IDENTITY, None)
                outs.append([dev.get(q_, shp) for q_, shp in ((py, (B * d.out_h * d.out_w * zcs,)), (gm, (Cout,)), (gv, (Cout,)), (sm_, (Cout,)), (ss_, (Cout,)))])
            for a_, b_ in zip(*outs):
                assert rel_err(b_, a_) <= 2e-5
        # db through the entry point itself: column-sum pass (option off), ones column of the GEMM (option on, TF32-exact delta)
        from oracle.net_oracle import tf32_round
        dq = tf32_round(delta)
        pdq = dev.put(to_nhwc(dq, zcs))
        dbq = dq.sum(axis=(0, 2, 3), dtype=np.float64)
        for opt in (0, 1):
            _lib.call("dl4jb200_ctx_set_option", ctx.handle, _lib.OPT_TF32_ACTIVATIONS, opt)
            pdw2, pdb2 = dev.empty(w.nbytes), dev.empty(Cout * 4)
            _lib.call("dl4jb200_conv_rowpack_wgrad", ctx.handle, C.byref(d), pxp, pitch, pdq, zcs, pdw2, pdb2)
            assert rel_err(dev.get(pdb2, (Cout,)), dbq) <= 1e-5, opt
    finally:
        _lib.call("dl4jb200_ctx_set_option", ctx.handle, _lib.OPT_TF32_ACTIVATIONS, 0)
        dev.free()
