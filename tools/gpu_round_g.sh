#!/bin/bash
# Round G: bench lines for BASELINE configs 2, 4, 5 (one GPU).
set -u
mkdir -p gpurun_out
for w in bulk10s longform60s bulk30s; do
  timeout 900 python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  tail -2 gpurun_out/bench_$w.err
  python -c "
import json,sys
d=json.load(open('gpurun_out/bench_$w.json'))
print({k:d[k] for k in ('value','ms_per_step','clocks')}, d['e2e']['value'], d['roofline']['frac'], d['path_frac_of_tensor_roofline'], d.get('cpu_baseline'))
"
done
