"""Whole-step TF32 error survey: runs the small parity networks in TF32 mode with and without
DL4JB200_OPT_TF32_ACTIVATIONS and prints the worst per-block gradient / parameter / score deviations from the CPU
oracle (no assertions; the numbers justify the tolerances in tests/test_gpu_network.py)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import test_gpu_network as T  # noqa: E402
from deeplearning4j_b200 import _lib, network, zoo  # noqa: E402
from oracle.net_oracle import OracleNet  # noqa: E402


def run(model, B, K, algo, rnd, fast, emulate=False):
    P, layout = zoo.param_layout(model)
    p0 = zoo.init_params(model)
    x, lab = T._data(model, B, K)
    net = network.Network(model, B, algo=algo, params=p0, round_tf32=rnd)
    ref = OracleNet(model, layout, P, p0, fast=fast, tf32=emulate, unrounded=net.fused_bn.keys())
    s_ref = ref.fit(x, lab)
    s = net.fit(x, lab)
    net.sync()
    s = float(s.cpu()[0])
    g, p = net.grads.cpu().numpy(), net.params.cpu().numpy()
    ge = T._block_errs(layout, g, ref.grads, 1e-3 * float(np.abs(ref.grads).max()))
    pe = T._block_errs(layout, p, ref.params, 1e-3 * float(np.abs(ref.params).max()), T._noise_blocks(layout, ref.grads))
    net.close()
    return abs(s - s_ref) / max(1.0, abs(s_ref)), max(ge.values()), max(pe.values())


cases = [("lenet", zoo.lenet(), 16, 10, False), ("tiny_resnet", T._tiny_resnet(), 8, 10, False),
         ("vgg32", zoo.vgg16(num_labels=10, input_shape=(3, 32, 32)), 4, 10, True),
         ("alexnet", None, 0, 0, True)]
for name, m, B, K, fast in cases:
    if m is None:
        continue
    for label, algo, rnd in (("fp32", _lib.ALGO_FP32, False), ("tf32 trunc", _lib.ALGO_TC_TF32, False),
                             ("tf32 rounded", _lib.ALGO_TC_TF32, True)):
        es, eg, ep = run(m, B, K, algo, rnd, fast)
        if rnd:
            print(f"{name:12s} {label:13s} score {es:.2e}  grad {eg:.2e}  param {ep:.2e}", flush=True)
            es, eg, ep = run(m, B, K, algo, rnd, fast, emulate=True)
            label = "tf32 vs q-orc"
        print(f"{name:12s} {label:13s} score {es:.2e}  grad {eg:.2e}  param {ep:.2e}", flush=True)
