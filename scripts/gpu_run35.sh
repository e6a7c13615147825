mkdir -p gpurun_out
python - <<'PY'
import sys; sys.path.insert(0,'.'); sys.argv=['x','alexnet','gpurun_out/timeline_alexnet.json']
import json, os, torch
from torch.profiler import ProfilerActivity, profile
from deeplearning4j_b200 import network, zoo
import bench
model='alexnet'
B, K = bench.CONFIGS[model][0], bench.CONFIGS[model][1]
m = zoo.alexnet(dropout=0.5)
net = network.Network(m, B)
x, lab = bench.synthetic(m, B, K)
net.set_dropout_masks(bench.make_masks(m, net.shapes, B))
net.set_batch(torch.from_numpy(x).pin_memory(), torch.from_numpy(lab).pin_memory())
for _ in range(5): net.fit()
net.sync()
net.capture(include_input=False)
step = lambda: (net._relayout_input(), net.fit_graph())
for _ in range(3): step()
net.sync()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3): step()
    net.sync()
out='gpurun_out/timeline_alexnet.json'
prof.export_chrome_trace(out + ".trace.json")
tr = json.load(open(out + ".trace.json"))
ks = [t for t in tr["traceEvents"] if t.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")]
ks.sort(key=lambda t: t["ts"])
rows = [{"ts": t["ts"], "dur": t["dur"], "name": t["name"][:70], "stream": t.get("args", {}).get("stream"), "cat": t["cat"]} for t in ks]
json.dump(rows, open(out, "w")); os.remove(out + ".trace.json")
PY
python scripts/timeline_report.py gpurun_out/timeline_alexnet.json 3 2>/dev/null | head -36 | cut -c1-140
