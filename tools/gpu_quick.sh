#!/bin/bash
# Quick check: parity suite + per-layer table (+ unfused table for comparison when Q3TTS_BOTH=1).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/profile_layers.py 64 125 fp16 > gpurun_out/layers_fp16.txt 2>&1; echo "layers rc=$?"
sed -n '1,2p;62,200p' gpurun_out/layers_fp16.txt
if [ "${Q3TTS_BOTH:-0}" = "1" ]; then
  Q3TTS_NO_FUSE=1 timeout 300 python tools/profile_layers.py 64 125 fp16 > gpurun_out/layers_fp16_nofuse.txt 2>&1; echo "layers nofuse rc=$?"
  sed -n '1,2p;72,200p' gpurun_out/layers_fp16_nofuse.txt
fi
