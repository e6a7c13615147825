"""Prints the last kernels of one step on both streams (from scripts/timeline.py output): where the step's tail goes."""
import json, collections, sys
rows = json.load(open(sys.argv[1]))
by = collections.defaultdict(list)
for r in rows:
    by[r['stream']].append(r)
streams = sorted(by.items(), key=lambda kv: -len(kv[1]))
main, side = streams[0][1], streams[1][1]
idx = [i for i, r in enumerate(main) if 'k_nchw_to_rowpack4' in r['name'] or 'k_nchw_to_nhwc' in r['name']]
t_end = main[idx[-1]]['ts']
span = float(sys.argv[2]) if len(sys.argv) > 2 else 1500.0
for nm, st in (("main", main), ("side", side)):
    print(nm, "tail:")
    for r in st:
        if t_end - span < r['ts'] <= t_end + 150:
            print("  %8.1f +%7.1f  %s" % (r['ts'] - t_end, r['dur'], r['name'][:72]))
