"""Short target for ncu: N warm forwards of the bench workload (64 x 10 s) then exit.
    python tools/ncu_target.py [B] [T] [passes] [mode]
One forward = 99 launches of which 71 are tcgen05 GEMM launches (see DESIGN.md "Launch list")."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 125
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 2
mode = sys.argv[4] if len(sys.argv) > 4 else "fp16"
cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model(compute_mode=mode)
dec = model.decoder
codes = make_codes(B, T, cfg, seed=1).cuda().transpose(1, 2).contiguous()
for _ in range(passes):
    wav = dec(codes)
torch.cuda.synchronize()
print("ok", tuple(wav.shape), dec.launch_count())
