"""dx of one convolution (FP32 and TF32 algos) against the oracle -- diagnostic. args: B Cin H W Cout k"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gpu_util import rel_err
from deeplearning4j_b200 import helpers as H
from oracle import oracle
B, Cin, Hh, W, Cout, k = (int(a) for a in sys.argv[1:7])
ctx = H.Context(0)
rng = np.random.default_rng(1)
x = rng.random((B, Cin, Hh, W), dtype=np.float32)
w = (rng.standard_normal((Cout, Cin, k, k)) * 0.1).astype(np.float32)
delta = rng.standard_normal((B, Cout, Hh, W)).astype(np.float32)
dwr, dbr, dxr = oracle.conv_bwd(x, w, delta, stride=(1, 1), pad=(0, 0), dilation=(1, 1), mode=oracle.SAME)
for name, algo in (("fp32", H.ALGO_FP32), ("tf32", H.ALGO_TC_TF32)):
    helper = H.CudnnConvolutionHelper(ctx, algo=algo)
    flat = np.zeros(Cout + w.size, np.float32)
    bview, wview = flat[:Cout].reshape(1, Cout), flat[Cout:].reshape(w.shape)
    grad, dx = helper.backpropGradient(x, w, delta, (k, k), (1, 1), (0, 0), bview, wview, "identity", None, None, None, "Same", (1, 1))
    err = np.abs(dx - dxr)
    print(name, "dW", rel_err(wview, dwr), "dx", rel_err(dx, dxr), "worst dx at", np.unravel_index(err.argmax(), err.shape),
          "per-channel max err (first 8 of sorted):", np.sort(err.max(axis=(0, 2, 3)))[-4:], "channels", np.argsort(err.max(axis=(0, 2, 3)))[-4:])
