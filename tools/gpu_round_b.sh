#!/bin/bash
# Per-layer table, ncu launch list of the bench command, ncu --set full of every GEMM launch of one forward
# (exported to CSV on the box: the .ncu-rep with sources is too large to bring back), small source-level reports of 5 launches.
set -u
mkdir -p gpurun_out
timeout 300 python tools/profile_layers.py 64 125 fp16 > gpurun_out/layers_fp16.txt 2>&1; echo "layers rc=$?"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 300 python tools/ncu_target.py 64 125 2 > gpurun_out/target.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none -k regex:"conv_gemm_tc_kernel|ws_conv_kernel" -s 71 -c 71 \
    -o /tmp/prof_all -f python tools/ncu_target.py 64 125 2 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/prof_all.ncu-rep --page raw --csv > gpurun_out/gemm_all_raw.csv 2> gpurun_out/ncu_export.err; echo "export rc=$?"
for idx in 43 57 58 64 65; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:"conv_gemm_tc_kernel|ws_conv_kernel" -s $((71+idx)) -c 1 \
      -o gpurun_out/prof_gemm_$idx -f python tools/ncu_target.py 64 125 2 > gpurun_out/ncu_k$idx.log 2>&1; echo "ncu $idx rc=$?"
done
find gpurun_out -size +20M -print -delete
du -sh gpurun_out; ls -la gpurun_out
cat gpurun_out/layers_fp16.txt
