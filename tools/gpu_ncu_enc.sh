#!/bin/bash
# `ncu --set full` of the two Listener recurrence kernels (tools/profile_enc.py), after the same command ran plain.
set -u
mkdir -p gpurun_out
timeout 300 python tools/profile_enc.py > gpurun_out/plain_enc.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_enc.log; exit 1; }
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k regex:'lstm_cl' -f -o gpurun_out/prof_enc python tools/profile_enc.py > gpurun_out/ncu_enc.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_enc.log
