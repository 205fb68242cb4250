#!/bin/bash
timeout 600 python -m pytest tests/test_attention_gpu.py tests/test_ctc_gpu.py -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/kernel_bench.py attn 2>&1 | grep "fwd\|bwd"
python tools/attn_trace.py 2>&1 | tail -14
