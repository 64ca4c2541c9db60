#!/bin/bash
# ncu --set full of (a) the o_proj GEMM of transformer layer 0 (general kernel, fp32 residual stream epilogue), (b) the transposed conv 192 -> 96.
set -u
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py 64 125 1 > gpurun_out/target.log 2>&1 || { echo target failed; tail -5 gpurun_out/target.log; exit 1; }
# conv_gemm_tc_kernel launches of one forward: 0 in_proj, then per layer qkv, o_proj, gate/up, down -> o_proj of layer 0 is index 2
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_tc_kernel -s 2 -c 1 -o gpurun_out/prof_oproj -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_oproj.log 2>&1; echo "ncu oproj rc=$?"
python tools/ncu_sass_summary.py gpurun_out/prof_oproj.ncu-rep 0.2 > gpurun_out/ncu_oproj.txt 2>&1
head -60 gpurun_out/ncu_oproj.txt
# slab_conv_kernel launches: pre_conv, dec0, tconv1, 3x conv7(768), tconv2, 3x conv7(384), tconv3, tconv4 -> tconv 192->96 is index 11
timeout 600 ncu --set full --clock-control none --import-source on -k regex:slab_conv_kernel -s 11 -c 1 -o gpurun_out/prof_tconv96 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_tconv96.log 2>&1; echo "ncu tconv96 rc=$?"
python tools/ncu_sass_summary.py gpurun_out/prof_tconv96.ncu-rep 8 > gpurun_out/ncu_tconv96.txt 2>&1
head -60 gpurun_out/ncu_tconv96.txt
