#!/bin/bash
# Round J: GPU tests, per-op table + per-stage roofline, default bench line, the three workloads, ncu launch list of the bench command.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu.log
PROFILE_REPS=4 timeout 300 python tools/profile_layers.py 64 125 > gpurun_out/layers_bulk.txt 2>&1; echo "layers rc=$?"
python tools/roofline_table.py gpurun_out/layers_bulk.txt > gpurun_out/roofline_by_stage.txt 2>&1; tail -30 gpurun_out/roofline_by_stage.txt
timeout 900 python bench.py > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
python -c "
import json
d=json.load(open('gpurun_out/bench_full.json'))
print({k:d[k] for k in ('value','ms_per_step','clocks','gpu_launches')}, 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'roof', d['roofline']['frac'], 'path', d['path_frac_of_tensor_roofline'])
print(d.get('cpu_baseline')); print([ (s['frames_per_packet'], s['device_ms']['p50'], s['e2e_ms']['p50'], s['e2e_ms']['p99']) for s in d['streaming']])
"
for w in longform60s bulk30s; do
  timeout 900 python bench.py --workload $w --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"
  python -c "
import json
d=json.load(open('gpurun_out/bench_$w.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['clocks']['sm_mhz'])"
done
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
