"""Eager vs CUDA-graph step time for ResNet-50, and a check that both produce the same parameters."""
import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
from deeplearning4j_b200 import network, zoo
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
m = zoo.resnet50()
p0 = zoo.init_params(m)
rng = np.random.default_rng(12345)
x = rng.random((B, 3, 224, 224), dtype=np.float32); lab = np.zeros((B, 1000), np.float32); lab[:, 1] = 1
def run(graph):
    net = network.Network(m, B, params=p0)
    net.fit(x, lab); net.fit()   # 2 eager warm-up steps
    if graph: net.capture()
    step = net.fit_graph if graph else net.fit
    for _ in range(2): step()
    net.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(net.stream)
    for _ in range(10): s = step()
    e1.record(net.stream); net.sync()
    ms = e0.elapsed_time(e1) / 10
    p = net.params.cpu().numpy().copy(); sc = float(s.cpu()[0])
    net.close(); del net; torch.cuda.empty_cache()
    return ms, p, sc
ms_e, p_e, s_e = run(False)
ms_g, p_g, s_g = run(True)
print(f"eager {ms_e:.3f} ms/step ({B/ms_e*1e3:.0f} img/s)  graph {ms_g:.3f} ms/step ({B/ms_g*1e3:.0f} img/s)")
print("params identical:", np.array_equal(p_e, p_g), "max diff", float(np.abs(p_e - p_g).max()), "scores", s_e, s_g)
