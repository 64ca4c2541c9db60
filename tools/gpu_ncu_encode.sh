#!/bin/bash
# ncu evidence for the encode path (run under gpurun, one GPU): launch list of the bench command (share of the step per kernel) and
# --set full of its dominant kernel (slab_conv_kernel, the stage-0 strided conv) and of the activation + split pass.
set -u
mkdir -p gpurun_out
B="python bench.py --workload encode10s --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $B > gpurun_out/r2e_bench_plain.json 2> gpurun_out/r2e_bench_plain.err || { echo "bench failed"; tail -3 gpurun_out/r2e_bench_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r2e_launches.csv $B > gpurun_out/r2e_ncu_launches.log 2>&1; echo "launch list rc=$?"
T="python tools/encode_profile.py 16 10"
for spec in "slab_conv_kernel:0:enc_slab_stage0" "ws_conv_kernel:1:enc_ws_k1" "k_enc_attention_mma:0:enc_attention_mma" "conv_gemm_tc_kernel:0:enc_gemm_k1_stage1"; do
  IFS=: read -r pat skip name <<< "$spec"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -o gpurun_out/r2e_$name -f $T > gpurun_out/r2e_ncu_$name.log 2>&1; echo "$name rc=$?"
  python tools/ncu_sass_summary.py gpurun_out/r2e_$name.ncu-rep 8 > gpurun_out/r2e_ncu_$name.txt 2>&1
  head -11 gpurun_out/r2e_ncu_$name.txt
done
