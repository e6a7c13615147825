"""What-if timings of the ResNet-50 step graph with parts of the work removed (numerically invalid; only to see how much
of the step each part accounts for once everything overlaps): usage whatif.py [nowgrad] [nobnbwd] [nofwdbranch]"""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
from deeplearning4j_b200 import network, zoo
B = 256
m = zoo.resnet50()
rng = np.random.default_rng(12345)
x = rng.random((B, 3, 224, 224), dtype=np.float32); lab = np.zeros((B, 1000), np.float32); lab[:, 1] = 1
def run(flags):
    net = network.Network(m, B)
    net.fit(x, lab); net.fit()
    if "nowgrad" in flags:
        orig = net._call_side
        net._call_side = lambda fn, *a: None if fn in ("dl4jb200_conv_bwd", "dl4jb200_conv_rowpack_wgrad") else orig(fn, *a)
    if "nofwdbranch" in flags:
        net.fwd_side = set()
    if "nobnbwd" in flags:
        orig2 = net._call
        def c2(fn, *a):
            if fn == "dl4jb200_bn_bwd":
                return None
            return orig2(fn, *a)
        net._call = c2
    net.capture()
    for _ in range(2): net.fit_graph()
    net.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(net.stream)
    for _ in range(10): net.fit_graph()
    e1.record(net.stream); net.sync()
    ms = e0.elapsed_time(e1) / 10
    net.close(); del net; torch.cuda.empty_cache()
    return ms
for flags in ([], ["nowgrad"], ["nofwdbranch"]):
    print(flags, f"{run(flags):.3f} ms/step", flush=True)
