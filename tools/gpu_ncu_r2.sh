#!/bin/bash
# Round-2 ncu evidence (run under gpurun, one GPU).  (1) launch list of the bench command (per-launch times: cold-cache, serialised -
# the kernels' SHARE of the step is what must agree with bench.py's in-step events); (2) --set full of the bandwidth kernels VERDICT
# r1 asked for (k_rvq, k_rmsnorm, k_dwconv_ln) and of the dominant kernel (slab_conv_kernel: conv7 C = 384, launch 7 of a forward).
set -u
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-streaming --no-cpu-baseline --no-torch-gpu"
timeout 300 $B > gpurun_out/r2n_bench_plain.json 2> gpurun_out/r2n_bench_plain.err || { echo "bench failed"; tail -3 gpurun_out/r2n_bench_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r2n_launches.csv $B > gpurun_out/r2n_ncu_launches.log 2>&1; echo "launch list rc=$?"
T="python tools/ncu_target.py 64 125 1"
timeout 300 $T > gpurun_out/r2n_target.log 2>&1 || { echo target failed; exit 1; }
for spec in "k_rvq:0:rvq" "k_rmsnorm:3:rmsnorm" "k_dwconv_ln:1:dwconv_ln" "k_split_f32:3:split_f32" "slab_conv_kernel:7:slab_conv7_c384"; do
  IFS=: read -r pat skip name <<< "$spec"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$pat -s $skip -c 1 -o gpurun_out/r2n_$name -f $T > gpurun_out/r2n_ncu_$name.log 2>&1; echo "$name rc=$?"
  python tools/ncu_sass_summary.py gpurun_out/r2n_$name.ncu-rep 8 > gpurun_out/r2n_ncu_$name.txt 2>&1
  head -11 gpurun_out/r2n_ncu_$name.txt
done
