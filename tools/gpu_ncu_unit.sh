#!/bin/bash
# ncu --set full with source of one fused-unit launch (C = 96) and one 1x1 conv launch (C = 384).
set -u
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:unit_kernel -s 3 -c 1 -o gpurun_out/prof_unit96 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_unit96.log 2>&1; echo "ncu unit96 rc=$?"
tail -3 gpurun_out/ncu_unit96.log
ls -la gpurun_out/*.ncu-rep
