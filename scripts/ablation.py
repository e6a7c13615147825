"""Same-box ablation of the step-level optimisations: ResNet-50 batch 256, CUDA-graph replay, one switch turned off at a
time (all variants are numerically equivalent training steps). Prints a markdown table (profiles/r01_ablation.md)."""
import sys
sys.path.insert(0, ".")
import numpy as np, torch
from deeplearning4j_b200 import network, zoo

B = 256
m = zoo.resnet50()
rng = np.random.default_rng(12345)
x = rng.random((B, 3, 224, 224), dtype=np.float32); lab = np.zeros((B, 1000), np.float32); lab[:, 1] = 1


def run(setup, graph=True):
    net = network.Network(m, B)
    setup(net)
    net.fit(x, lab); net.fit()
    if graph:
        net.capture()
    step = net.fit_graph if graph else net.fit
    for _ in range(3):
        step()
    net.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(net.stream)
    for _ in range(20):
        step()
    e1.record(net.stream); net.sync()
    ms = e0.elapsed_time(e1) / 20
    net.close(); del net; torch.cuda.empty_cache()
    return ms


def no_prepack(net):
    net.prepack = False
    net._call("dl4jb200_filters_invalidate")


variants = [
    ("everything on", lambda n: None, True),
    ("no CUDA graph (eager launches)", lambda n: None, False),
    ("filter gradients / shortcut branches on the main stream", lambda n: setattr(n, "overlap_wgrad", False), True),
    ("no epsilon fan-in + ReLU fusion in the backprop-data epilogue", lambda n: setattr(n, "fuse_dgrad", False), True),
    ("no batch-norm statistics in the convolution epilogue", lambda n: setattr(n, "fuse_bn_stats", False), True),
    ("filters packed per call instead of once per step", no_prepack, True),
]
base = None
print("| variant | ms/step | images/s | vs everything on |")
print("|---|---:|---:|---:|")
for name, setup, graph in variants:
    ms = run(setup, graph)
    base = base or ms
    print(f"| {name} | {ms:.2f} | {B / ms * 1e3:.0f} | {ms / base - 1:+.1%} |", flush=True)
