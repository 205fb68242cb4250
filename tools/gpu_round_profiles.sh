#!/bin/bash
# Round-end evidence: stand-alone kernel timings, phase timeline, GPU-busy split, and the ncu launch list of one train step.
set -u
mkdir -p gpurun_out
timeout 600 python tools/kernel_bench.py all > gpurun_out/kernel_bench.txt 2>&1; echo "kernel_bench rc=$?"
timeout 300 python tools/attn_trace.py > gpurun_out/attn_phases.txt 2>&1; echo "trace rc=$?"
timeout 600 python tools/step_busy.py 2>&1 | grep -v Warn > gpurun_out/step_busy.txt; echo "busy rc=$?"
python tools/profile_step.py 2 > gpurun_out/plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python tools/profile_step.py 2 > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
