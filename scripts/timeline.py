"""Kernel timeline of the captured training step as it really runs (CUPTI through torch.profiler): per-kernel start / end on
both streams of a graph replay -- what ncu's serialised, cold-cache launch list cannot show (overlap, gaps, L2-warm times).
Usage: python scripts/timeline.py [model] [out.json]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile
from deeplearning4j_b200 import network, zoo
import bench

model = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/timeline.json"
graph = os.environ.get("TL_EAGER", "0") != "1"
B, K = bench.CONFIGS[model][0], bench.CONFIGS[model][1]
m = zoo.MODELS[model]()
net = network.Network(m, B)
x, lab = bench.synthetic(m, B, K)
net.set_batch(torch.from_numpy(x).pin_memory(), torch.from_numpy(lab).pin_memory())
for _ in range(5):
    net.fit()
net.sync()
if graph:
    net.capture(include_input=False)
    step = lambda: (net._relayout_input(), net.fit_graph())
else:
    step = lambda: (net._relayout_input(), net.fit())
for _ in range(3):
    step()
net.sync()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step()
    net.sync()
ev = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None:
        ev.append((e.time_range.start, e.time_range.end - e.time_range.start, e.name, getattr(e, "device_resource_id", None) if hasattr(e, "device_resource_id") else None))
# fall back to the chrome trace for stream ids
prof.export_chrome_trace(out + ".trace.json")
tr = json.load(open(out + ".trace.json"))
ks = [t for t in tr["traceEvents"] if t.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ks.sort(key=lambda t: t["ts"])
rows = [{"ts": t["ts"], "dur": t["dur"], "name": t["name"][:70], "stream": t.get("args", {}).get("stream"), "cat": t["cat"]} for t in ks]
json.dump(rows, open(out, "w"))
os.remove(out + ".trace.json")
print(len(rows), "device events over 3 steps")
