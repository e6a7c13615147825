"""Where does the end-to-end loop of bench.py lose time against the device-resident loop? (diagnostic)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from deeplearning4j_b200 import network, zoo
import bench

model = sys.argv[1] if len(sys.argv) > 1 else "vgg16"
B, K = bench.CONFIGS[model][0], bench.CONFIGS[model][1]
m = zoo.MODELS[model]()
net = network.Network(m, B)
x, lab = bench.synthetic(m, B, K)
xh, labh = torch.from_numpy(x).pin_memory(), torch.from_numpy(lab).pin_memory()
net.set_batch(xh, labh)
for _ in range(5): net.fit()
net.sync()
net.capture(include_input=False)
def timed(name, fn, n=10):
    net.sync()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    t0 = time.perf_counter()
    ev[0].record(net.stream)
    for i in range(n):
        fn()
        ev[i + 1].record(net.stream)
    net.sync()
    wall = (time.perf_counter() - t0) / n * 1e3
    print(name, "wall %.2f ms/step" % wall, "gpu", [round(ev[i].elapsed_time(ev[i + 1]), 2) for i in range(n)])
def g():
    net._relayout_input(); net.fit_graph()
def gs():
    net._relayout_input(); net.fit_graph(); net.sync()
def gsc():
    net._relayout_input(); s = net.fit_graph(); net.sync(); float(s.cpu()[0])
timed("graph back-to-back", g)
timed("graph + sync", gs)
timed("graph + sync + score read", gsc)
net.prefetch(xh, labh)
def e2e():
    net.prefetch(xh, labh)
    s = net.fit_prefetched(); net.sync(); float(s.cpu()[0])
timed("prefetch e2e", e2e)
def e2e_nosync():
    net.prefetch(xh, labh)
    net.fit_prefetched()
timed("prefetch, no per-step sync", e2e_nosync)
