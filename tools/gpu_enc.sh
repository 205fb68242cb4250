#!/bin/bash
# Encoder recurrence kernels: parity tests, stand-alone timings, then the whole suite + bench.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_encoder_gpu.py -x -q > gpurun_out/pytest_enc.log 2>&1; echo "enc pytest rc=$?"
tail -25 gpurun_out/pytest_enc.log
timeout 300 python tools/enc_bench.py > gpurun_out/enc_bench.txt 2>&1; echo "enc bench rc=$?"
tail -8 gpurun_out/enc_bench.txt
if [ "${1:-}" = "full" ]; then
  timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
  tail -5 gpurun_out/pytest_gpu.log
  timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; echo "bench rc=$?"
  tail -1 gpurun_out/bench.log | cut -c1-600; tail -3 gpurun_out/bench.err
fi
