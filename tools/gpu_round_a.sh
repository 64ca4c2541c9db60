#!/bin/bash
# First GPU pass of the session: parity tests, bench line, per-op table, ncu launch list + full capture of every GEMM launch.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python tools/profile_layers.py 64 125 fp16 > gpurun_out/layers_fp16.txt 2>&1; echo "layers rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python tools/ncu_target.py 64 125 2 > gpurun_out/target.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"conv_gemm_tc_kernel|ws_conv_kernel" -s 71 -c 71 \
    -o gpurun_out/prof_gemm_r1 -f python tools/ncu_target.py 64 125 2 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json; tail -5 gpurun_out/ncu_full.log
