#!/bin/bash
timeout 900 python -m pytest tests/test_ctc_gpu.py -m gpu -x -q 2>&1 | tail -6
timeout 300 python tools/exp_ctc_steps.py 2>&1 | tail -8
bash tools/gpu_ctc_times.sh
