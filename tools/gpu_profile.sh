#!/bin/bash
# ncu launch list of one whole train step + full captures of our kernels.  Run via gpurun.
set -u
mkdir -p gpurun_out
python tools/profile_step.py 2 > gpurun_out/plain.log 2>&1 &&
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches.csv python tools/profile_step.py 2 > gpurun_out/ncu_list.log 2>&1
echo "list rc=$?"
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:'attn_step|ctc_' -s 40 -c 6 \
    -f -o gpurun_out/prof_kernels python tools/profile_step.py 2 > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"
ls -la gpurun_out
