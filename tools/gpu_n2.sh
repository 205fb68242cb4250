#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.log 2> gpurun_out/bench_n2.err; echo "n2 rc=$?"
tail -1 gpurun_out/bench_n2.log | cut -c1-400; tail -3 gpurun_out/bench_n2.err
