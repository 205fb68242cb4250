"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.  usage: launch_summary.py launches.csv [top]"""
import csv, sys
from collections import defaultdict
rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
h = rows[0]
kn, mv = h.index("Kernel Name"), h.index("Metric Value")
agg = defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[mv].replace(",", ""))
    except ValueError:
        continue
    a = agg[r[kn][:100]]; a[0] += 1; a[1] += v / 1e3
tot = sum(a[1] for a in agg.values()); n = sum(a[0] for a in agg.values())
print("# total launches %d, total %.1f us (per-launch times are cold-cache and serialised: compare SHARES)" % (n, tot))
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[: int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print("%6d launches %10.1f us %5.1f%%  %s" % (c, t, 100 * t / tot, k))
