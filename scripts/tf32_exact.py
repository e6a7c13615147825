"""Does the TF32 tensor-core path reproduce an fp32-accumulated convolution when its operands are TF32-representable?
(diagnostic behind DL4JB200_OPT_TF32_ACTIVATIONS and OracleNet(tf32=True))"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from gpu_util import rel_err
from deeplearning4j_b200 import helpers as H
from oracle import oracle
from oracle.net_oracle import tf32_round as q, tf32_trunc as t

ctx = H.Context(0)
rng = np.random.default_rng(1)
for B, Cin, Hh, W, Cout, k, s, p in ((16, 1, 28, 28, 20, (5, 5), (1, 1), (0, 0)), (4, 64, 14, 14, 64, (3, 3), (1, 1), (1, 1)),
                                     (4, 256, 14, 14, 64, (1, 1), (1, 1), (0, 0))):
    x = rng.random((B, Cin, Hh, W), dtype=np.float32)
    w = (rng.standard_normal((Cout, Cin) + k) * 0.5).astype(np.float32)
    b = rng.standard_normal(Cout).astype(np.float32)
    helper = H.CudnnConvolutionHelper(ctx, algo=H.ALGO_TC_TF32)
    for label, xx, ww in (("raw operands vs raw oracle", x, w), ("rounded operands, same oracle", q(x), q(w)),
                          ("raw on GPU vs oracle(trunc x, rna w)", x, w)):
        z = helper.preOutput(xx, ww, b, k, s, p, None, None, "Truncate", (1, 1))
        if label.startswith("raw on GPU"):
            zr = oracle.conv_fwd(t(x), q(w), b, stride=s, pad=p, dilation=(1, 1), mode=oracle.TRUNCATE)
        else:
            zr = oracle.conv_fwd(xx, ww, b, stride=s, pad=p, dilation=(1, 1), mode=oracle.TRUNCATE)
        delta = q(rng.standard_normal(zr.shape).astype(np.float32)) if "rounded" in label else rng.standard_normal(zr.shape).astype(np.float32)
        flat = np.zeros(Cout + w.size, np.float32)
        bview, wview = flat[:Cout].reshape(1, Cout), flat[Cout:].reshape(w.shape)
        grad, dx = helper.backpropGradient(xx, ww, delta, k, s, p, bview, wview, "identity", None, None, None, "Truncate", (1, 1))
        if label.startswith("raw on GPU"):
            dwr, dbr, dxr = oracle.conv_bwd(t(x), q(w), t(delta), stride=s, pad=p, dilation=(1, 1), mode=oracle.TRUNCATE)
        else:
            dwr, dbr, dxr = oracle.conv_bwd(xx, ww, delta, stride=s, pad=p, dilation=(1, 1), mode=oracle.TRUNCATE)
        print(f"{(B, Cin, Hh, W, Cout, k)} {label:40s} z {rel_err(z, zr):.2e} dW {rel_err(wview, dwr):.2e} dx {rel_err(dx, dxr):.2e}", flush=True)
