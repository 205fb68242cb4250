#!/bin/bash
set -u
mkdir -p gpurun_out
for cfg in "" "B2ASR_ATT_CS=8" "B2ASR_ATT_CS=8 B2ASR_ATT_CAP_KB=113" "B2ASR_ATT_CS=2" "B2ASR_ATT_CAP_KB=113"; do
  echo "== $cfg"
  env $cfg timeout 200 python tools/kernel_bench.py attn 2>&1 | tail -4
done
