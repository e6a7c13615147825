"""Per-convolution lower bounds for the ResNet-50 step: max(flops / TF32 peak, algorithmic bytes / HBM peak) per call,
next to the measured per-call times of a bench.py --dump-profile file. Shows which layers are tensor-bound and which
are HBM-bound with fp32-resident activations, and how far each is from its own bound."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deeplearning4j_b200 import zoo  # noqa: E402

prof = json.load(open(sys.argv[1]))["per_layer"]
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
model = sys.argv[3] if len(sys.argv) > 3 else "resnet50"
pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
TF, BW = pk["bf16_tflops_sustained"] / 2 * 1e12, pk["hbm_gbs"] * 1e9
try:  # the measured TF32 dense peak of this pool (scripts/measure_tf32_peak.py), else the provisional half-BF16 figure
    TF = json.load(open(os.path.join(ROOT, "profiles", "r02_tf32_peak.json")))["tf32_tflops_sustained"] * 1e12
except Exception:  # noqa: BLE001
    pass
from oracle import oracle as _o  # noqa: E402  (shape arithmetic without the CUDA library: runs on the CPU box)
m = zoo.MODELS[model]()
shp = zoo.infer_shapes(m, _o.out_size)
vm = {v["name"]: v for v in m["vertices"]}
rows = []
for v in m["vertices"]:
    if v["op"] not in ("conv", "dense", "output"):
        continue
    n = v["name"]
    so, si = shp[n], shp[v["in"][0]]
    if v["op"] == "conv":
        kh, kw = v["kernel"]
        nout, nin = B * so[0] * so[1] * so[2], B * si[0] * si[1] * si[2]
        fl = 2.0 * B * so[1] * so[2] * so[0] * si[0] * kh * kw
    else:
        cin = int(si[0] * (si[1] if len(si) > 1 else 1) * (si[2] if len(si) > 2 else 1))
        nout, nin = B * so[0], B * cin
        fl = 2.0 * B * so[0] * cin
    first = vm[v["in"][0]]["op"] in ("input", "zeropad")
    for ps, key, f, by in (("fwd", "conv_fwd", fl, 4.0 * (nin + nout)),
                           ("bwd", "conv_bwd", fl * (1 if first else 2), 4.0 * (nin + nout + (0 if first else nin + nout)))):
        meas = None
        for k2, e in prof.items():
            tag, fn = k2.rsplit(":", 1)
            if tag == n and ((ps == "fwd" and fn.endswith("fwd")) or (ps == "bwd" and not fn.endswith("fwd"))):
                meas = (meas or 0) + e["ms"]
        tc, tm = f / TF * 1e3, by / BW * 1e3
        rows.append((n, ps, f / 1e9, by / 1e6, tc, tm, meas))
tot_b = tot_m = tot_c = tot_h = 0.0
print(f"{'layer':22s} pass  GFLOP     MB   t_tc   t_hbm  bound   meas   meas/bound")
for n, ps, gf, mb, tc, tm, meas in rows:
    bd = max(tc, tm)
    if meas is None:
        continue
    tot_b += bd; tot_m += meas; tot_c += tc; tot_h += tm
    print(f"{n:22s} {ps:4s} {gf:7.1f} {mb:7.1f} {tc:6.3f} {tm:6.3f} {'tc ' if tc > tm else 'hbm'} {meas:7.3f} {meas / bd:6.2f}")
print(f"total: bound {tot_b:.2f} ms (pure tensor {tot_c:.2f}, pure hbm {tot_h:.2f}); measured {tot_m:.2f} ms; ratio {tot_m / tot_b:.2f}")
