#!/bin/bash
for cfg in "0 40960 227" "8 20480 112" "8 16384 112" "8 12288 112"; do
  set -- $cfg
  echo "== CS=$1 STAGE=$2 CAP=$3"
  B2ASR_ATT_CS=$1 B2ASR_ATT_STAGE=$2 B2ASR_ATT_CAP_KB=$3 timeout 300 python tools/kernel_bench.py attn 2>&1 | grep "fwd\|bwd" | grep same
done
B2ASR_ATT_CS=8 B2ASR_ATT_STAGE=16384 B2ASR_ATT_CAP_KB=112 python tools/attn_trace.py 2>&1 | tail -28
