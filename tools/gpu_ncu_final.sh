#!/bin/bash
# ncu --set full of the shipped C = 96 unit (4 + 12 warps), the C = 192 unit (phase 1 on 16 warps) and the transposed conv 192 -> 96 (12 epilogue warps).
set -u
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py 64 125 1 > gpurun_out/target.log 2>&1 || { echo target failed; tail -5 gpurun_out/target.log; exit 1; }
# unit_kernel launches of one forward: 0-2 = C 192 (dil 1, 3, 9), 3-5 = C 96
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unit_kernel -s 3 -c 1 -o gpurun_out/prof_unit96_p2x -f python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_a.log 2>&1; echo "unit96 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unit_kernel -s 0 -c 1 -o gpurun_out/prof_unit192_p1a -f python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_b.log 2>&1; echo "unit192 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:slab_conv_kernel -s 11 -c 1 -o gpurun_out/prof_tconv96_12w -f python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_c.log 2>&1; echo "tconv96 rc=$?"
for n in unit96_p2x unit192_p1a tconv96_12w; do python tools/ncu_sass_summary.py gpurun_out/prof_$n.ncu-rep 8 > gpurun_out/ncu_$n.txt 2>&1; head -12 gpurun_out/ncu_$n.txt; grep stall_ gpurun_out/ncu_$n.txt | tr "\n" " "; echo; done
