"""Decoder-loop forward / backward ms per batch and the train step, from bench.py's per-call CUDA-event pass (M2)."""
import json, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-decode", "--no-cpu", "--no-extra", "--steps", "6"],
                     capture_output=True, text=True)
line = [l for l in out.stdout.splitlines() if l.startswith("{")]
if not line:
    print(out.stdout[-2000:], out.stderr[-3000:]); sys.exit(1)
d = json.loads(line[-1])
k = d["kernels"]
print("train step %.2f ms (%.1f utt/s), e2e %.1f utt/s" % (d["ms_per_step"], d["value"], d["e2e"]["value"]))
for n in sorted(k, key=lambda n: -k[n]["ms_per_step"]):
    print("  %-28s %7.3f ms/step  %8.1f us/call x %.0f" % (n, k[n]["ms_per_step"], k[n]["avg_us"], k[n]["calls_per_step"]))
