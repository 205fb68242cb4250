#!/bin/bash
# A/B of an environment knob on the train-step bench: tools/gpu_ab.sh VAR  (runs VAR=0 then VAR=1)
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_decoder_gpu.py tests/test_trainstep_gpu.py tests/test_encoder_gpu.py -x -q 2>&1 | tail -2
for v in 0 1; do
  env $1=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu > gpurun_out/bench_$1_$v.log 2> gpurun_out/bench_$1_$v.err
  echo "$1=$v rc=$? $(python -c "import json,sys; d=json.loads(open('gpurun_out/bench_$1_$v.log').read().strip().splitlines()[-1]); print(d['ms_per_step'], {k:round(v['avg_us'],1) for k,v in d['kernels'].items() if 'loop' in k or 'lstm' in k})")"
done
