"""Aggregate ncu warp-stall samples per CUDA source line.
usage: python tools/ncu_lines.py report.ncu-rep <kernel-id e.g. :::1 or regex name> [top]"""
import csv
import subprocess
import sys
from collections import defaultdict

rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
sel = ["--kernel-id", kid] if kid.startswith(":") else ["-k", "regex:" + kid, "-c", "1"]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"] + sel,
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
agg = defaultdict(lambda: [0, 0, ""])       # (file,line) -> samples, instrs, text
fname = ""
cur_line = ""
hdr = None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Line No":
        hdr = r
        si = hdr.index("# Samples")
        ii = hdr.index("Instructions Executed")
        continue
    if hdr is None or len(r) <= si:
        continue
    line, text = r[0], r[1]
    if line == "":
        line, text = cur_line, ""
    else:
        cur_line = line
    try:
        s = int(r[si] or 0)
        n = int(r[ii] or 0)
    except ValueError:
        continue
    key = (fname, line)
    agg[key][0] += s
    agg[key][1] += n
    if text:
        agg[key][2] = text
tot = sum(v[0] for v in agg.values()) or 1
toti = sum(v[1] for v in agg.values()) or 1
print("total samples %d, warp instructions %d" % (tot, toti))
for (f, l), (s, n, t) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%5.1f%% smp %5.1f%% ins  %s:%s  %s" % (100.0 * s / tot, 100.0 * n / toti, f, l, t.strip()[:110]))
