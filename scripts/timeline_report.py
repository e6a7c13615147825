"""Summarises scripts/timeline.py output: step time, per-stream busy time, gaps on the main stream, top kernels."""
import json, sys, collections
rows = json.load(open(sys.argv[1]))
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
t0, t1 = rows[0]["ts"], max(r["ts"] + r["dur"] for r in rows)
print("span %.3f ms for %d steps -> %.3f ms/step; %d events/step" % ((t1 - t0) / 1e3, nsteps, (t1 - t0) / 1e3 / nsteps, len(rows) // nsteps))
by_stream = collections.defaultdict(list)
for r in rows:
    by_stream[r["stream"]].append(r)
for s, rs in sorted(by_stream.items(), key=lambda kv: -len(kv[1])):
    busy = sum(r["dur"] for r in rs)
    gaps = [rs[i + 1]["ts"] - (rs[i]["ts"] + rs[i]["dur"]) for i in range(len(rs) - 1)]
    pos = [g for g in gaps if g > 0]
    print("stream %s: %d events, busy %.3f ms/step, positive gaps %.3f ms/step (%d gaps, median %.2f us, >5us: %d, >20us: %d)" % (
        s, len(rs) // nsteps, busy / 1e3 / nsteps, sum(pos) / 1e3 / nsteps, len(pos) // nsteps, sorted(pos)[len(pos) // 2] if pos else 0,
        sum(1 for g in pos if g > 5) // nsteps, sum(1 for g in pos if g > 20) // nsteps))
# union busy time (any stream)
iv = sorted((r["ts"], r["ts"] + r["dur"]) for r in rows)
u, cs, ce = 0.0, iv[0][0], iv[0][1]
for a, b in iv[1:]:
    if a > ce:
        u += ce - cs; cs, ce = a, b
    else:
        ce = max(ce, b)
u += ce - cs
print("GPU busy (union of all streams) %.3f ms/step, idle %.3f ms/step" % (u / 1e3 / nsteps, ((t1 - t0) - u) / 1e3 / nsteps))
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    a = agg[(r["name"], r["stream"])]
    a[0] += 1; a[1] += r["dur"]
print("top kernels (per step):")
for (n, s), (c, d) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:32]:
    print("  %-72s s=%-3s n=%4d  %8.1f us  avg %6.1f us" % (n, s, c // nsteps, d / nsteps, d / c))
# largest gaps on the busiest stream, with the neighbours
main = max(by_stream.values(), key=len)
g = sorted(((main[i + 1]["ts"] - (main[i]["ts"] + main[i]["dur"]), main[i]["name"][:40], main[i + 1]["name"][:40]) for i in range(len(main) - 1)), reverse=True)
print("largest main-stream gaps:")
for gap, a, b in g[:15]:
    print("  %7.1f us  after %-40s before %s" % (gap, a, b))
