#!/bin/bash
# One `ncu --set full` capture of the hot-path kernels at their section-8d shapes (tools/profile_kernels.py).
set -u
mkdir -p gpurun_out
export B2ASR_CTC_OVERLAP=0
timeout 300 python tools/profile_kernels.py > gpurun_out/plain_kernels.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_kernels.log; exit 1; }
timeout 1200 ncu --profile-from-start off --set full --clock-control none --import-source on \
    -k regex:'attn_keys2|gemm_nt|ctc_|attn_step' -f -o gpurun_out/prof_kernels python tools/profile_kernels.py > gpurun_out/ncu_kernels.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_kernels.log
