"""Per-vertex activation / per-block raw gradient deviation for full-size AlexNet (FP32 mode) -- diagnostic."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import test_gpu_network as T
from deeplearning4j_b200 import _lib, network, zoo
from oracle.net_oracle import OracleNet
m = zoo.alexnet()
B = 4
P, layout = zoo.param_layout(m)
p0 = zoo.init_params(m)
x, lab = T._data(m, B, 1000)
net = network.Network(m, B, algo=_lib.ALGO_FP32, params=p0)
if "nofuse" in sys.argv:
    net.fuse_dgrad = False
if "noprepack" in sys.argv:
    net.prepack = False
    net._call("dl4jb200_filters_invalidate")
if "nooverlap" in sys.argv:
    net.overlap_wgrad = False
ref = OracleNet(m, layout, P, p0, fast=False)
net.set_batch(x, lab); net.forward(True); net.backward(); net.sync()
ref.forward(x, lab, True); ref.backward()
for v in m["vertices"]:
    n = v["name"]
    if v["op"] in ("input", "zeropad"): continue
    got = net.output_host(n); r = ref.act[n].reshape(got.shape)
    print(f"act  {n:10s} {np.abs(got - r).max() / max(np.abs(r).max(), 1e-30):.2e}")
g = net.grads.cpu().numpy()
for name, key, off, shp, pk in layout:
    k = int(np.prod(shp)); a, r = g[off:off + k], ref.grads[off:off + k]
    print(f"grad {name:10s} {key:3s} max|ref| {np.abs(r).max():.3e} err/blockmax {np.abs(a - r).max() / max(np.abs(r).max(), 1e-30):.2e}")
off, k, shp, _ = net.pidx[("cnn3", "b")]
a, r = g[off:off + k], ref.grads[off:off + k]
err = np.abs(a - r)
print("cnn3.b worst channels", np.argsort(err)[-8:], err[np.argsort(err)[-8:]], "ref there", r[np.argsort(err)[-8:]])
print("channels with err > 1e-6:", int((err > 1e-6).sum()), "of", k)
e3 = net.eps["cnn3"].cpu().numpy()  # delta of cnn3, NHWC
print("delta3 shape", e3.shape, "nonzero frac", float((e3 != 0).mean()))
# oracle delta3 / delta4 recomputed step by step from the oracle's own activations
from oracle import oracle as o
def conv_dx(name, d):
    v = ref.vmap[name]
    u = v["in"][0]
    return o.conv_bwd(ref.lin[name], ref._p(name, "W"), d, stride=v["stride"], pad=v["pad"], dilation=v["dilation"],
                      mode={"Strict": o.STRICT, "Truncate": o.TRUNCATE, "Same": o.SAME}[v["mode"]])[2]
# walk back from the loss exactly as OracleNet.backward does, keeping the deltas
eps = {"ffn2": None}
d_out = ref.delta
dx = o.dense_bwd(ref.lin["output"], ref._p("output", "W"), d_out)[2]
d_f2 = o.act_bwd(ref.act["ffn2"], dx, o.ACT_RELU)
dx = o.dense_bwd(ref.lin["ffn2"], ref._p("ffn2", "W"), d_f2)[2]
d_f1 = o.act_bwd(ref.act["ffn1"], dx, o.ACT_RELU)
dx = o.dense_bwd(ref.lin["ffn1"], ref._p("ffn1", "W"), d_f1)[2].reshape(ref.act["maxpool3"].shape)
v = ref.vmap["maxpool3"]
dx = o.pool_bwd(ref.lin["maxpool3"], dx, v["kernel"], v["stride"], v["pad"], (1, 1), o.SAME, o.POOL_MAX)
d5 = o.act_bwd(ref.act["cnn5"], dx, o.ACT_RELU)
d4 = o.act_bwd(ref.act["cnn4"], conv_dx("cnn5", d5), o.ACT_RELU)
e3_ref = conv_dx("cnn4", d4)
d3 = o.act_bwd(ref.act["cnn3"], e3_ref, o.ACT_RELU)
got3 = np.ascontiguousarray(e3[..., :384].transpose(0, 3, 1, 2))
got4 = np.ascontiguousarray(net.eps["cnn4"].cpu().numpy()[..., :256].transpose(0, 3, 1, 2))
print("delta4 err", np.abs(got4 - d4).max() / np.abs(d4).max(), " delta3 err", np.abs(got3 - d3).max() / np.abs(d3).max())
bad = np.argwhere(np.abs(got3 - d3) > 1e-3 * np.abs(d3).max())
print("bad delta3 elements", len(bad), "of", d3.size, "first", bad[:10].tolist())
for b_, c_, y_, x_ in bad[:6]:
    print("  at", (b_, c_, y_, x_), "got", got3[b_, c_, y_, x_], "ref", d3[b_, c_, y_, x_], "eps_ref", e3_ref[b_, c_, y_, x_], "a3 ref", ref.act["cnn3"][b_, c_, y_, x_],
          "a3 gpu", net.output_host("cnn3")[b_, c_, y_, x_])
