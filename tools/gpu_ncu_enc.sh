#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/enc_one.py 2 > gpurun_out/enc_one.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/enc_one.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lstm_cl -c 2 -f -o gpurun_out/enc_cluster \
    python tools/enc_one.py 1 > gpurun_out/ncu_enc.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/ncu_enc.log
ls -la gpurun_out/*.ncu-rep
