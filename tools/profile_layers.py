"""Per-op table of one bulk-decode step: ms, achieved TFLOP/s and minimum-traffic GB/s per launch.
    python tools/profile_layers.py [B] [T] [mode]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 125
mode = sys.argv[3] if len(sys.argv) > 3 else "fp16"
cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model(compute_mode=mode)
dec = model.decoder
codes = make_codes(B, T, cfg, seed=1).cuda().transpose(1, 2).contiguous()
for _ in range(3):
    dec(codes)
reps = int(os.environ.get("PROFILE_REPS", "1"))  # average over several profiled forwards (same-box A/B: cuts the run-to-run noise)
dec.set_profile(True)
acc = None
for _ in range(reps):
    dec(codes)
    pr = dec.get_profile()
    acc = [list(x) for x in pr] if acc is None else [[a[0], a[1] + b[1], a[2], a[3], a[4]] for a, b in zip(acc, pr)]
prof = [(a[0], a[1] / reps, a[2], a[3], a[4]) for a in acc]
dec.set_profile(False)
names = {0: "rvq", 1: "gemm", 2: "rmsnorm", 3: "attn", 4: "dwln", 5: "tail", 6: "unit", 7: "wsconv", 8: "slabconv", 9: "sattn", 10: "state", 11: "split"}
tot = sum(p[1] for p in prof)
print(f"B={B} T={T} mode={mode} total {tot:.3f} ms over {len(prof)} launches")
print(f"{'#':>3} {'op':8} {'rows':>9} {'N':>5} {'Cin':>5} {'k':>2} {'ms':>8} {'%':>5} {'TFLOP/s':>8} {'minGB/s':>8}")
for i, (ty, ms, fl, by, d) in enumerate(prof):
    print(f"{i:3d} {names[ty]:8} {d[0]:9d} {d[1]:5d} {d[2]:5d} {d[3]:2d} {ms:8.4f} {100*ms/tot:5.1f} {fl/ms/1e9 if ms > 0 else 0:8.1f} {by/ms/1e6 if ms > 0 else 0:8.1f}")
