#!/bin/bash
set -u
mkdir -p gpurun_out
python tools/profile_decode.py > gpurun_out/plain_dec.log 2>&1 &&
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:beam_step -c 2 \
    -f -o gpurun_out/prof_beam python tools/profile_decode.py > gpurun_out/ncu_beam.log 2>&1
echo "beam rc=$?"
python tools/profile_step.py 2 > gpurun_out/plain.log 2>&1 &&
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:attn_step_bwd -s 20 -c 2 \
    -f -o gpurun_out/prof_attn_bwd python tools/profile_step.py 2 > gpurun_out/ncu_bwd.log 2>&1
echo "bwd rc=$?"
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:ctc_ -c 2 \
    -f -o gpurun_out/prof_ctc python tools/profile_step.py 2 > gpurun_out/ncu_ctc.log 2>&1
echo "ctc rc=$?"
