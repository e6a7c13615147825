"""-m gpu: the CTA-pair form of the persistent convolution GEMM (DL4JB200_OPT_PAIR_TILES: 256-row tiles, tcgen05 cta_group::2,
each CTA loading half of the filter tile). Off by default (measured slower on ResNet-50's shapes, profiles/r02_pair_tiles.md) but
kept correct: convolution outputs are bit-identical with the option on, and the convolution / fused-epilogue / epilogue-statistics
/ strided-backprop suites pass unchanged when every context of the process defaults to it."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from gpu_util import Dev, to_nhwc

pytestmark = pytest.mark.gpu

SHAPES = [  # B, Cin, H, W, Cout, k, stride, mode
    (8, 64, 28, 28, 64, 3, 1, "Same"),       # 64-wide tiles, 49 row tiles: an odd count, the last pair is half empty
    (8, 128, 14, 14, 512, 1, 1, "Truncate"),  # 256-wide tiles, B-resident filter
    (8, 512, 14, 14, 128, 1, 1, "Truncate"),  # 128-wide tiles, 16 K-chunks
    (3, 256, 7, 7, 256, 3, 1, "Same"),        # ragged: 147 rows = one full and one partial 128-row half
    (8, 256, 14, 14, 128, 1, 2, "Truncate"),  # strided 1x1: the backprop-data scatters rows (REMAP epilogue)
]


@pytest.mark.parametrize("shape", SHAPES)
def test_pair_tiles_bit_identical(shape):
    from deeplearning4j_b200 import _lib as L, helpers as H
    B, Cin, Hh, W, Cout, k, s, mode = shape
    rng = np.random.default_rng(2)
    x = rng.standard_normal((B, Cin, Hh, W)).astype(np.float32)
    w = (rng.standard_normal((Cout, Cin, k, k)) * 0.1).astype(np.float32)
    b = rng.standard_normal(Cout).astype(np.float32)
    d = H.conv_desc(x.shape, Cout, (k, k), (s, s), (0, 0), (1, 1), mode)
    dz = rng.standard_normal((B, Cout, d.out_h, d.out_w)).astype(np.float32)
    ctx = H.Context(0)
    L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_TF32_ACTIVATIONS, 1)
    dev = Dev(ctx)
    try:
        px, pw, pb, pdz = dev.put(to_nhwc(x)), dev.put(w), dev.put(b), dev.put(to_nhwc(dz))
        n_in, n_out = x.size, dz.size
        res = []
        for pair in (0, 1):
            L.call("dl4jb200_ctx_set_option", ctx.handle, L.OPT_PAIR_TILES, pair)
            pz, pdx = dev.empty(n_out * 4, fill=0xFF), dev.empty(n_in * 4, fill=0xFF)
            L.call("dl4jb200_conv_fwd", ctx.handle, C.byref(d), px, L.NHWC, Cin, pw, pb, pz, L.NHWC, Cout, L.ACT_RELU, L.ALGO_TC_TF32)
            L.call("dl4jb200_conv_bwd_data_fused", ctx.handle, C.byref(d), pw, pdz, Cout, pdx, Cin, None, None, L.ACT_IDENTITY, L.ALGO_TC_TF32)
            res.append((dev.get(pz, (n_out,)), dev.get(pdx, (n_in,))))
        np.testing.assert_array_equal(res[0][0], res[1][0])
        np.testing.assert_array_equal(res[0][1], res[1][1])
        assert np.abs(res[1][0]).max() > 0 and np.isfinite(res[1][1]).all()
    finally:
        dev.free()
        ctx.close()


def test_suites_pass_with_pair_tiles_as_default():
    """Re-runs the convolution-facing suites in a child process whose contexts default to pair tiles (DL4JB200_PAIR_TILES=1):
    oracle parity, fused epilogues, forward statistics / backward sums in the epilogue, strided backprop, whole-step replay."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, DL4JB200_PAIR_TILES="1")
    files = ["tests/test_gpu_helpers.py", "tests/test_gpu_bwd_bnstats.py", "tests/test_gpu_strided_bwd.py", "tests/test_gpu_bn_maxpool.py",
             "tests/test_gpu_replay.py"]
    r = subprocess.run([sys.executable, "-m", "pytest", "-m", "gpu", "-q", "-x", "-k", "conv or bwd or strided or maxpool or replay"] + files,
                       cwd=root, env=env, capture_output=True, text=True, timeout=1500)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-1000:]
