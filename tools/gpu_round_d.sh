#!/bin/bash
# Round D: parity suite + per-layer table + short bench with the fused unit kernel in the plan.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python tools/profile_layers.py 64 125 fp16 > gpurun_out/layers_fp16.txt 2>&1; echo "layers rc=$?"
sed -n '1,2p;62,200p' gpurun_out/layers_fp16.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench rc=$?"
cat gpurun_out/bench_short.json; tail -5 gpurun_out/bench_short.err
