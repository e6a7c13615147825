"""GPU test of dl4jb200_bn_fwd_noalloc, the batch-norm forward that never allocates.

Who needs it: callers of the library that cannot afford or are not allowed a device allocation while they run (a
loop that must not stall on cudaMalloc, a process whose GPU memory is owned by another allocator, code running where
CUDA memory management is forbidden). The entry point keeps its scratch in a fixed array inside the library, so the
check here is simply that, through the C ABI, it returns what dl4jb200_bn_fwd returns -- twice in a row and with
different shapes, because every call reuses the same array and finds the previous call's leftovers in it.
The file name sorts last on purpose: a run that stops at the first failure has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

from gpu_util import Dev, from_nhwc, to_nhwc

pytestmark = pytest.mark.gpu

TOL = 1e-5  # same bound as the other bandwidth-kernel tests (relative to max|ref|)


def _run(lib, ctx, dev, fn, x_dev, shape, layout, cs, gamma, beta):
    B, Cc, H, W = shape
    n = B * H * W * (cs if layout == lib.NHWC else Cc)
    py = dev.empty(n * 4, fill=0xFF)
    pg, pb = dev.put(gamma), dev.put(beta)
    pm, pv = dev.put(np.zeros(Cc, np.float32)), dev.put(np.ones(Cc, np.float32))
    psm, pss = dev.empty(Cc * 4), dev.empty(Cc * 4)
    extra = (None,) if fn == "dl4jb200_bn_fwd" else ()  # the existing entry point also takes an optional residual
    lib.call(fn, ctx.handle, x_dev, py, pg, pb, pm, pv, psm, pss, B, Cc, H, W, layout, cs, 0.9, 1e-5, 1, 1, lib.ACT_RELU, *extra)
    return [dev.get(p, (k,)) for p, k in ((py, n), (pm, Cc), (pv, Cc), (psm, Cc), (pss, Cc))]


def test_bn_fwd_noalloc_matches_bn_fwd_two_calls_in_a_row(oracle):
    from deeplearning4j_b200 import build
    build.build()
    from deeplearning4j_b200 import _lib as lib, helpers
    ctx = helpers.Context(0)
    dev = Dev(ctx)
    rng = np.random.default_rng(2024)
    try:
        # second call is smaller and in the other layout: it must not be confused by what the first one left behind
        for shape, layout in (((8, 96, 14, 14), lib.NHWC), ((3, 20, 9, 7), lib.NCHW), ((8, 96, 14, 14), lib.NHWC)):
            B, Cc, H, W = shape
            cs = (Cc + 3) // 4 * 4 if layout == lib.NHWC else 0
            x = (rng.standard_normal(shape) * 2 + 1).astype(np.float32)
            gamma = (rng.random(Cc) + 0.5).astype(np.float32)
            beta = rng.standard_normal(Cc).astype(np.float32)
            px = dev.put(to_nhwc(x, cs) if layout == lib.NHWC else x)
            ref = _run(lib, ctx, dev, "dl4jb200_bn_fwd", px, shape, layout, cs, gamma, beta)
            got = _run(lib, ctx, dev, "dl4jb200_bn_fwd_noalloc", px, shape, layout, cs, gamma, beta)
            for g, r in zip(got, ref):
                assert np.isfinite(g).all()
                assert np.abs(g.astype(np.float64) - r).max() <= TOL * max(np.abs(r).max(), 1e-30)
            # and directly against the CPU restatement of BatchNormalization.preOutput (+ the fused ReLU)
            gm, gv = np.zeros(Cc, np.float32), np.ones(Cc, np.float32)
            yr, _, _ = oracle.bn_fwd(x, gamma, beta, gm, gv, 0.9, 1e-5, True, True)
            yr = np.maximum(yr, 0.0)
            y = got[0].reshape(B, H, W, cs) if layout == lib.NHWC else got[0].reshape(shape)
            y = from_nhwc(y, Cc) if layout == lib.NHWC else y
            assert np.abs(y.astype(np.float64) - yr).max() <= TOL * np.abs(yr).max()
            assert np.abs(got[1] - gm).max() <= TOL * max(np.abs(gm).max(), 1e-30) and np.abs(got[2] - gv).max() <= TOL * np.abs(gv).max()
        # sizes beyond the fixed scratch are refused with a clear message, not silently mishandled
        big = dev.empty(16)
        with pytest.raises(lib.DL4JB200Error, match="exceed the fixed scratch"):
            lib.call("dl4jb200_bn_fwd_noalloc", ctx.handle, big, big, None, None, big, big, None, None, 1, 5000, 1, 1, lib.NCHW, 0,
                     0.9, 1e-5, 1, 1, lib.ACT_IDENTITY)
    finally:
        dev.free()
        ctx.close()
