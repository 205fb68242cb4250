"""BASELINE.json config 5 (CTC stress sweep) timed alone: bench.ctc_m5_leg with the overlapped and the serial schedule.
    python tools/ctc_m5.py [iters]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dev = torch.device("cuda:0")
for mode in ("2", "3", "0"):
    os.environ["B2ASR_CTC_OVERLAP"] = mode
    r = bench.ctc_m5_leg(dev, iters=iters)
    print("B2ASR_CTC_OVERLAP=%s (peak %.0f GB/s %s)" % (mode, r["peak_gbs"], r["peak_source"]))
    for p in r["points"]:
        print("  T'=%4d L=%3d  %8.1f us  %6.3f GB  %7.1f GB/s  frac %.3f  ws %.0f MB finite=%s" % (
            p["frames"], p["label_len"], p["us"], p["alg_gb"], p["achieved_gbs"], p["frac"], p["workspace_mb"], p["finite"]))
    if mode == "2":
        json.dump(r, open(os.path.join(ROOT, "gpurun_out", "ctc_m5.json"), "w"))
