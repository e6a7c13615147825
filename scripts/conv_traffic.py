"""profiles/r02_conv_traffic.json from the one-step ncu launch list (agg_launches.py --out): DRAM bytes and time of every
convolution-family launch (k_conv_tc_* / k_wgrad_finalize* / k_stem_wgrad_finalize / k_pack_*). usage: conv_traffic.py step.csv out.json"""
import collections, csv, json, sys
rows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
L = collections.OrderedDict()
for r in rows:
    e = L.setdefault(r["ID"], {"name": r["Kernel Name"]})
    e[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
fam = [e for e in L.values() if any(k in e["name"] for k in ("k_conv_tc_", "k_wgrad_finalize", "k_stem_wgrad_finalize", "k_pack_"))]
out = {"launches": len(fam), "dram_read_bytes": sum(e.get("dram__bytes_read.sum", 0.0) for e in fam),
       "dram_write_bytes": sum(e.get("dram__bytes_write.sum", 0.0) for e in fam),
       "kernel_us": round(sum(e.get("gpu__time_duration.sum", 0.0) for e in fam) / 1e3, 2),
       "source": "profiles/r02_launches_step.csv (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none; "
                 "one steady-state ResNet-50 step at batch 256): every k_conv_tc_* / k_wgrad_finalize* / k_pack_* launch. Since round 2 the "
                 "backprop-data epilogues also read the batch-norm input / ReLU mask / fan-in sum of the fused vertices (that traffic used to "
                 "belong to k_bn_partial_nhwc and the elementwise kernels)."}
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(out["launches"], "launches", round((out["dram_read_bytes"] + out["dram_write_bytes"]) / 1e9, 2), "GB", out["kernel_us"], "us")
