#!/bin/bash
for w in timit_m1 libri_m2s; do
  timeout 900 python bench.py --workload $w --steps 6 --warmup 3 --no-decode > gpurun_out/bench_$w.log 2> gpurun_out/bench_$w.err; echo "$w rc=$?"; tail -c 300 gpurun_out/bench_$w.err
done
