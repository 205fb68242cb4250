#!/bin/bash
# `ncu --set full` of three launches of the fused beam step at the M4 shape (tools/profile_decode.py), after the same
# command ran plain.
set -u
mkdir -p gpurun_out
timeout 300 python tools/profile_decode.py > gpurun_out/plain_beam.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_beam.log; exit 1; }
timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k regex:'beam_step' -c 3 -f -o gpurun_out/prof_beam python tools/profile_decode.py > gpurun_out/ncu_beam.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_beam.log
