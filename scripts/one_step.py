"""Runs W warm-up + K ResNet-50 training steps (no timing): the command profiled under ncu for profiles/."""
import sys
sys.path.insert(0, ".")
import numpy as np
import torch
from deeplearning4j_b200 import network, zoo, _lib
W, K = int(sys.argv[1]), int(sys.argv[2])
B = int(sys.argv[3]) if len(sys.argv) > 3 else 256
m = zoo.resnet50()
net = network.Network(m, B)
rng = np.random.default_rng(12345)
x = rng.random((B, 3, 224, 224), dtype=np.float32); lab = np.zeros((B, 1000), np.float32); lab[:, 1] = 1
net.set_batch(x, lab)
for _ in range(W): net.fit()
net.sync()
import time
n0 = _lib.cdll().dl4jb200_launch_count()
t0 = time.perf_counter()
for _ in range(K): net.fit()
t1 = time.perf_counter()
net.sync()
t2 = time.perf_counter()
print(f"host issue time {1e3*(t1-t0)/max(K,1):.2f} ms/step, wall {1e3*(t2-t0)/max(K,1):.2f} ms/step")
print("launches per step", (_lib.cdll().dl4jb200_launch_count() - n0) / max(K, 1), "warm launches", n0)
