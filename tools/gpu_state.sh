#!/bin/bash
# Current-state evidence: GPU tests, smoke, bench line, step busy split.
set -u
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
tail -2 gpurun_out/bench.log; tail -3 gpurun_out/bench.err
timeout 600 python tools/step_busy.py 2>&1 | grep -v Warn > gpurun_out/step_busy.txt; echo "busy rc=$?"
