#!/bin/bash
# Stream-step latency with and without CUDA graphs (1 and 256 streams), after the GPU test suite.
Q3TTS_STREAM_GRAPH=1 timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -3
for S in 1 256; do for E in Q3TTS_X=0 Q3TTS_STREAM_GRAPH=1; do echo -n "$S $E "; env $E timeout 300 python bench.py --stream-only --streams $S 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print([(s['frames_per_packet'], round(s['device_ms']['p50'],3), round(s['e2e_ms']['p50'],3), round(s['e2e_ms']['p99'],3)) for s in d['streaming']])"; done; done
