#!/bin/bash
# Round C: GPU parity suite on the current tree, fused-unit selftest, per-layer table, short bench.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/unit_selftest.py > gpurun_out/unit_selftest.log 2>&1; echo "unit selftest rc=$?"
cat gpurun_out/unit_selftest.log
timeout 300 python tools/profile_layers.py 64 125 fp16 > gpurun_out/layers_fp16.txt 2>&1; echo "layers rc=$?"
sed -n '1,2p;62,200p' gpurun_out/layers_fp16.txt
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err; echo "bench rc=$?"
cat gpurun_out/bench_short.json
