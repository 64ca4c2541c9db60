#!/bin/bash
# ncu --set full with source of the last slab-stream launch of one forward (transposed conv 192 -> 96, Snake epilogue), single-CTA form.
set -u
mkdir -p gpurun_out
Q3TTS_NO_PAIR=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:slab_conv_kernel -s 10 -c 1 -o gpurun_out/prof_op88 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_op88.log 2>&1; echo "ncu op88 rc=$?"
tail -2 gpurun_out/ncu_op88.log
python tools/ncu_sass_summary.py gpurun_out/prof_op88.ncu-rep 8 > gpurun_out/ncu_op88.txt 2>&1
head -60 gpurun_out/ncu_op88.txt
