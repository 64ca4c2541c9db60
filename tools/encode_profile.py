"""GPU tool: per-op profile of the encode path (events around every launch) for B items of S seconds.
    python tools/encode_profile.py [B] [seconds]"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.synthetic import make_audio

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
sec = float(sys.argv[2]) if len(sys.argv) > 2 else 10.0
T = int(sec * 24000)
NAMES = {1: "gemm", 7: "ws_conv", 8: "slab_conv", 12: "conv0", 13: "act_split", 14: "layernorm", 15: "attention", 16: "latents_tap", 17: "split_cols", 18: "argmin_update"}
m = synthetic_model(device="cuda:0", with_encoder=True)
wav = make_audio(B, T, seed=0).cuda()
for _ in range(3):
    m.encode(wav)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5):
    m.encode(wav)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 5
print(f"encode {B} x {sec:g} s: {ms:.2f} ms per call = {B * sec / ms * 1e3:.0f} audio-s/s (device-resident wav)")
d = m.decoder
d.set_profile(True)
m.encode(wav)
prof = d.get_profile()
d.set_profile(False)
tot = sum(p[1] for p in prof)
print(f"profiled: {len(prof)} launches, {tot:.2f} ms")
agg = {}
for code, t, fl, by, dims in prof:
    k = (NAMES.get(code, str(code)), dims)
    e = agg.setdefault(k, [0, 0.0, 0.0, 0.0]); e[0] += 1; e[1] += t; e[2] += fl; e[3] += by
print(f"{'op':<11} {'rows':>9} {'n':>5} {'k':>6} {'taps':>4} {'launches':>8} {'ms':>8} {'%':>5} {'TFLOP/s':>8} {'GB/s':>7}")
for (name, dims), (n, t, fl, by) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:40]:
    print(f"{name:<11} {dims[0]:>9} {dims[1]:>5} {dims[2]:>6} {dims[3]:>4} {n:>8} {t:>8.3f} {100 * t / tot:>5.1f} {fl / t / 1e9 if t else 0:>8.1f} {by / t / 1e6 if t else 0:>7.0f}")
byname = {}
for (name, dims), (n, t, fl, by) in agg.items():
    e = byname.setdefault(name, [0, 0.0]); e[0] += n; e[1] += t
print({k: (v[0], round(v[1], 3)) for k, v in sorted(byname.items(), key=lambda kv: -kv[1][1])})
