#!/bin/bash
# Round H: per-op table, ncu launch list of the bench command, DRAM bytes per launch of one forward, full capture of the attention kernel.
set -u
mkdir -p gpurun_out
timeout 300 python tools/profile_layers.py 64 125 > gpurun_out/layers_bulk.txt 2>&1; echo "layers rc=$?"
python tools/roofline_table.py gpurun_out/layers_bulk.txt > gpurun_out/roofline_by_stage.txt 2>&1; tail -32 gpurun_out/roofline_by_stage.txt
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-streaming > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'conv_gemm_tc_kernel|slab_conv_kernel|unit_kernel|tail_conv_kernel|k_attention|k_rvq|k_rmsnorm|k_dwconv_ln|ws_conv_kernel' \
    -c 200 --csv --log-file gpurun_out/dram_fwd.csv python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_dram.log 2>&1; echo "ncu dram rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_attention -s 2 -c 1 -o gpurun_out/prof_attention -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_attn.log 2>&1; echo "ncu attn rc=$?"
python tools/ncu_sass_summary.py gpurun_out/prof_attention.ncu-rep 1 > gpurun_out/ncu_attention.txt 2>&1
head -45 gpurun_out/ncu_attention.txt
