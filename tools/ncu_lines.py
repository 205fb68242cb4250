"""Attribute an `ncu --page source --csv` dump (SASS rows) to CUDA source lines through `nvdisasm -g` of the object file.
    python tools/ncu_lines.py <report.ncu-rep> <object.o> <mangled-kernel-substring> [top]
Prints, per source line: warp instructions executed, stall samples, and the share of each."""
import csv, io, os, re, subprocess, sys, tempfile

rep, obj, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
sel = ["--kernel-name", "regex:" + os.environ["NCU_KERNEL"]] if os.environ.get("NCU_KERNEL") else []   # multi-kernel reports
raw = subprocess.run(["ncu", "-i", rep] + sel + ["--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ci = {n: hdr.index(n) for n in ("Source", "# Samples", "Instructions Executed")}
sass = [(r[ci["Source"]].strip(), int(r[ci["# Samples"]] or 0), int(r[ci["Instructions Executed"]] or 0)) for r in rows[hdr_i + 1:] if len(r) > ci["Instructions Executed"]]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and kern in l)
lines, cur = [], None
for l in dis[start + 1:]:
    if l.startswith("//-----") or (l.startswith(".text.") and kern not in l):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", l)
    if m:
        lines.append(cur)
print("SASS rows in report %d, disassembled instructions %d" % (len(sass), len(lines)))
n = min(len(sass), len(lines))
agg = {}
for (src, smp, ins), loc in zip(sass[:n], lines[:n]):
    a = agg.setdefault(loc, [0, 0])
    a[0] += ins; a[1] += smp
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print("total warp instructions %d, samples %d" % (ti, ts))
srcs = {}
for (f, ln), (ins, smp) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    if f not in srcs:
        for root in ("end-to-end-asr-pytorch_b200/csrc",):
            pth = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), root, f)
            srcs[f] = open(pth).read().splitlines() if os.path.exists(pth) else []
    text = srcs[f][ln - 1].strip()[:90] if srcs[f] and ln - 1 < len(srcs[f]) else ""
    print("%-14s %4d  inst %5.1f%%  samples %5.1f%%  %s" % (f, ln, 100.0 * ins / max(ti, 1), 100.0 * smp / max(ts, 1), text))
