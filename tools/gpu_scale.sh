#!/bin/bash
# tools/gpu_scale.sh N : bench line at N GPUs of one box (utterance-sharded, no data-path collective); last stdout line = the JSON line
set -u
N=$1
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/bench_${N}gpu.err | tail -1 > gpurun_out/bench_${N}gpu.json
echo "rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_${N}gpu.json')); print(d['n_gpus'], round(d['value']), round(d['ms_per_step'],2), round(d['e2e']['value']), d['clocks'])"
