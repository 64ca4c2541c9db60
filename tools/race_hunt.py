"""Find the first stage whose probe is not bitwise reproducible.  python tools/race_hunt.py [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model()
dec = model.decoder
codes = make_codes(64, 125, cfg, seed=1).cuda().transpose(1, 2).contiguous()
stages = ["quantizer", "pre_conv", "pre_transformer", "upsample.0.0", "upsample.0.1", "upsample.1.0", "upsample.1.1", "decoder.0.snake"]
stages += [f"decoder.{i}.block.{j}" for i in range(1, 5) for j in (1, 2, 3, 4)]
for st in stages:
    ref = dec.probe(codes, st).clone()
    bad = 0
    where = None
    for r in range(reps):
        out = dec.probe(codes, st)
        if not torch.equal(out, ref):
            bad += 1
            if where is None:
                d = (out - ref).abs()
                idx = torch.nonzero(d > 0)
                where = (idx[0].tolist(), idx[-1].tolist(), idx.shape[0], d.max().item())
    print(f"{st:22s} shape {tuple(ref.shape)}  {bad}/{reps} differ  {where}", flush=True)
    if bad and "--all" not in sys.argv:
        break
