"""Per-block gradient deviation of the TF32 path from the TF32-quantised oracle (diagnostic)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_network as T
from deeplearning4j_b200 import _lib, network, zoo
from oracle.net_oracle import OracleNet

which = sys.argv[1] if len(sys.argv) > 1 else "lenet"
model, B, K, fast = {"lenet": (zoo.lenet(), 16, 10, False), "tiny": (T._tiny_resnet(), 8, 10, False),
                     "vgg": (zoo.vgg16(num_labels=10, input_shape=(3, 32, 32)), 4, 10, True)}[which]
P, layout = zoo.param_layout(model)
p0 = zoo.init_params(model)
x, lab = T._data(model, B, K)
net = network.Network(model, B, algo=_lib.ALGO_TC_TF32, params=p0)
ref = OracleNet(model, layout, P, p0, fast=fast, tf32=True, unrounded=net.fused_bn.keys())
net.set_batch(x, lab); net.forward(True); net.backward(); net.sync()
ref.forward(x, lab, True); ref.backward()
ref0 = OracleNet(model, layout, P, p0, fast=fast)
ref0.forward(x, lab, True)
inp = net.out[model["vertices"][0]["name"]].cpu().numpy()
print("input low mantissa bits set on device:", int((inp.view(np.uint32) & 0x1FFF).astype(bool).sum()), "of", inp.size,
      "round_tf32", net.round_tf32)
for v in model["vertices"]:
    n = v["name"]
    if v["op"] in ("input", "zeropad") or n in net.fused_bn: continue
    got = net.output_host(n)
    r = ref.act[n].reshape(got.shape)
    r0 = ref0.act[n].reshape(got.shape)
    low = int((got.view(np.uint32) & 0x1FFF).astype(bool).sum())
    print(f"act  {n:14s} vs q-oracle {np.abs(got - r).max() / max(np.abs(r).max(), 1e-30):.2e}  vs plain {np.abs(got - r0).max() / max(np.abs(r0).max(), 1e-30):.2e}  unrounded elems {low}/{got.size}")
g = net.grads.cpu().numpy()
gm = float(np.abs(ref.grads).max())
for name, key, off, shp, pk in layout:
    n = int(np.prod(shp))
    a, r = g[off:off + n], ref.grads[off:off + n]
    print(f"grad {name:14s} {key:6s} max|ref| {np.abs(r).max():.3e} err/blockmax {np.abs(a - r).max() / max(np.abs(r).max(), 1e-30):.2e}")
