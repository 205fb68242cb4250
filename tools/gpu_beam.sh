#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_beam_gpu.py -m gpu -x -q > gpurun_out/pytest_beam.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/pytest_beam.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
