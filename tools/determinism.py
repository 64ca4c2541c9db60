"""Bitwise determinism of the bulk decode: N runs of the same 64 x 10 s batch.  python tools/determinism.py [runs]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

runs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model()
dec = model.decoder
codes = make_codes(64, 125, cfg, seed=1).cuda().transpose(1, 2).contiguous()
ref = dec(codes).clone()
bad = 0
for r in range(runs):
    out = dec(codes)
    if not torch.equal(out, ref):
        d = (out - ref).abs()
        idx = torch.nonzero(d.view(64, -1) > 0)
        print(f"run {r}: {idx.shape[0]} samples differ, max {d.max().item():.3e}; first at item {idx[0,0].item()} sample {idx[0,1].item()}, last at item {idx[-1,0].item()} sample {idx[-1,1].item()}")
        bad += 1
print(f"{bad} of {runs} runs differ")
