#!/bin/bash
timeout 600 python -m pytest tests/test_ctc_gpu.py -m gpu -x -q 2>&1 | tail -4
python tools/kernel_bench.py ctc 2>&1 | tail -8
python tools/profile_ctc.py > gpurun_out/plain_ctc.log 2>&1 && ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:ctc_lattice -c 1 -f -o gpurun_out/prof_ctc_m5b python tools/profile_ctc.py > gpurun_out/ncu_ctc5.log 2>&1; echo rc=$?
