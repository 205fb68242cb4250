#!/bin/bash
# ncu launch list of one train step of the bench workload (per-launch gpu__time_duration), summarised by kernel.
set -u
mkdir -p gpurun_out
python tools/profile_step.py 2 > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python tools/profile_step.py 2 > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
python tools/launch_summary.py gpurun_out/launches.csv 40 > gpurun_out/launches_summary.txt 2>&1
head -30 gpurun_out/launches_summary.txt
