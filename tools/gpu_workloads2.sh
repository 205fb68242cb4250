#!/bin/bash
set -u
mkdir -p gpurun_out
for wl in timit_m1 libri_m2s; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --no-decode > gpurun_out/bench_$wl.log 2> gpurun_out/bench_$wl.err; echo "$wl rc=$?"
  python -c "import json; d=json.loads(open('gpurun_out/bench_$wl.log').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['cpu_baseline'])"
done
