#!/bin/bash
# Round E: bench line (with CPU baseline), reference arm, ncu launch list of the bench command, ncu --set full of the two top kernels.
set -u
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$?"
cat gpurun_out/bench_full.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python tools/ncu_target.py 64 125 1 > gpurun_out/target.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unit_kernel -s 3 -c 1 -o gpurun_out/prof_unit96 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_unit96.log 2>&1; echo "ncu unit96 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unit_kernel -s 0 -c 1 -o gpurun_out/prof_unit192 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_unit192.log 2>&1; echo "ncu unit192 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:slab_conv_kernel -s 8 -c 1 -o gpurun_out/prof_slab384 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_slab384.log 2>&1; echo "ncu slab rc=$?"
ls -la gpurun_out/*.ncu-rep
