#!/bin/bash
# Round I: GPU tests, default bench line (with CPU baseline and streaming), reference arm.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_full.json'))
print({k:d[k] for k in ('value','ms_per_step','clocks','gpu_launches')}, 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'roof', d['roofline']['frac'], d['roofline']['traffic'], 'path', d['path_frac_of_tensor_roofline'])
print(d['kernels']); print(d.get('cpu_baseline')); print([ (s['frames_per_packet'], s['device_ms']['p50'], s['e2e_ms']['p50'], s['e2e_ms']['p99']) for s in d['streaming']])
"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cat gpurun_out/bench_ref.json
