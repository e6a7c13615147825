"""Aggregates an ncu `--metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv` launch list
by kernel name. usage: agg_launches.py list.csv [--last-step MARKER_KERNEL] [--out step.csv]
--last-step keeps only the launches from the last occurrence of MARKER_KERNEL on (one steady-state step)."""
import collections, csv, re, sys
args = sys.argv[1:]
path = args[0]
marker = args[args.index("--last-step") + 1] if "--last-step" in args else None
out = args[args.index("--out") + 1] if "--out" in args else None
lines = [l for l in open(path) if not l.startswith('==')]
rows = list(csv.DictReader(lines))
launches = collections.OrderedDict()  # ID -> {name, stream, metrics}
for r in rows:
    e = launches.setdefault(r['ID'], {"name": r['Kernel Name'], "stream": r['Stream']})
    v = float(r['Metric Value'].replace(',', '')); u = r['Metric Unit']
    if r['Metric Name'].startswith('gpu__time'):
        v = v / 1e3 if u == 'ns' else (v * 1e3 if u == 'ms' else (v * 1e6 if u == 's' else v))
    else:
        v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    e[r['Metric Name']] = v
L = list(launches.values())
if marker:
    idx = [i for i, e in enumerate(L) if marker in e["name"]]
    L = L[idx[-1]:]
if out:
    ids = set()
    keep = [k for k, e in launches.items() if e in L]
    with open(out, "w") as f:
        f.write(lines[0])
        w = csv.DictWriter(f, fieldnames=rows[0].keys(), quoting=csv.QUOTE_ALL)
        ks = set(keep)
        for r in rows:
            if r['ID'] in ks:
                w.writerow(r)
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0]); tot = 0
per_stream = collections.defaultdict(lambda: [0, 0.0])
for e in L:
    name = re.sub(r'\(.*', '', e["name"])
    t = e.get('gpu__time_duration.sum', 0.0)
    a = agg[name]; a[0] += 1; a[1] += t; a[2] += e.get('dram__bytes_read.sum', 0.0); a[3] += e.get('dram__bytes_write.sum', 0.0)
    tot += t
    per_stream[e["stream"]][0] += 1; per_stream[e["stream"]][1] += t
print(f"total {tot:.1f} us over {len(L)} launches; per stream: " + ", ".join(f"{k}: {n} launches {t:.0f} us" for k, (n, t) in per_stream.items()))
for k, (n, t, br, bw) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    extra = f"  dram rd {br/1e6:9.1f} MB wr {bw/1e6:9.1f} MB  {(br+bw)/max(t,1e-9)/1e6:6.2f} TB/s" if (br or bw) else ""
    print(f"{k[:60]:60s} n={n:4d} {t:9.1f} us {100*t/tot:5.1f}%{extra}")
