#!/bin/bash
# One ncu --set full capture of the attention backward kernel inside a train step.
set -u
mkdir -p gpurun_out
python tools/profile_step.py 2 > gpurun_out/plain.log 2>&1 &&
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:attn_step_bwd -s 20 -c 2 \
    -f -o gpurun_out/prof_attn_bwd python tools/profile_step.py 2 > gpurun_out/ncu_bwd.log 2>&1
echo "bwd rc=$?"
