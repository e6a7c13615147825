import sys, numpy as np
sys.path.insert(0, ".")
from deeplearning4j_b200 import zoo, network, _lib
from oracle.net_oracle import OracleNet
which, algo = sys.argv[1], int(sys.argv[2])
fast = len(sys.argv) > 3 and sys.argv[3] == "fast"
if which == "vgg":
    m = zoo.vgg16(num_labels=10, input_shape=(3, 32, 32)); B, K = 4, 10
elif which == "lenet":
    m = zoo.lenet(); B, K = 16, 10
elif which == "tiny":
    sys.path.insert(0, "tests"); import test_gpu_network as t; m = t._tiny_resnet(); B, K = 8, 10
else:
    m = zoo.resnet50(); B, K = 4, 1000
P, layout = zoo.param_layout(m); p0 = zoo.init_params(m)
rng = np.random.default_rng(12345)
x = rng.random((B,) + tuple(m["input"]), dtype=np.float32); lab = np.zeros((B, K), np.float32); lab[:, 1] = 1
ref = OracleNet(m, layout, P, p0, fast=fast)
net = network.Network(m, B, algo=algo, params=p0)
ref.forward(x, lab, True); ref.backward()
net.set_batch(x, lab); net.forward(True); net.backward(); net.sync()
def re(a, b): return float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-30))
print("== forward")
for v in m["vertices"]:
    n = v["name"]
    if v["op"] in ("input", "zeropad") or n in net.fused_bn: continue
    got = net.output_host(n); r = ref.act[n]
    print(f"{n:24s} {v['op']:8s} {re(got, r.reshape(got.shape)):.3e}")
print("== grads (reverse)")
g = net.grads.cpu().numpy(); gmax = np.abs(ref.grads).max()
for name, key, off, shp, pk in reversed(layout):
    n = int(np.prod(shp)); a, b = g[off:off+n], ref.grads[off:off+n]
    print(f"{name:24s} {key:6s} rel {re(a,b):.3e}  abs/gmax {np.abs(a-b).max()/gmax:.3e}  refmax {np.abs(b).max():.3e}")
