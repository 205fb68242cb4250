import os, sys
sys.path.insert(0, "/root/repo")
import torch, bench
dev = torch.device("cuda:0")
for K in ("0", "4", "8"):
    os.environ["B2ASR_CTC_K"] = K
    r = bench.ctc_m5_leg(dev, iters=5)
    print("K=%s: " % K + "  ".join("(%d,%d) %.0f us %.3f" % (p["frames"], p["label_len"], p["us"], p["frac"]) for p in r["points"]))
