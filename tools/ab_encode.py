"""One sample for same-box A/B runs of the ENCODE path: median device time of 64 x 10 s and the per-kernel totals with a given build.
    python tools/ab_encode.py <path/to/libq3tts_decode.so> [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qwen3_tts_b200 import _capi

_capi.set_library_path(sys.argv[1])
import ctypes
_probe = ctypes.CDLL(sys.argv[1])
for _n in list(_capi.EXPORTS) + list(_capi.DEBUG_EXPORTS):
    if not hasattr(_probe, _n):
        _capi.EXPORTS.pop(_n, None); _capi.DEBUG_EXPORTS.pop(_n, None)
import torch

from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.synthetic import make_audio

reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
m = synthetic_model(device="cuda:0", with_encoder=True)
wav = make_audio(64, 240000, seed=0).cuda()
for _ in range(3):
    m.encode(wav)
ts = []
for _ in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); m.encode(wav); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
ts.sort()
d = m.decoder
d.set_profile(True); d.get_profile()
acc = {}
for _ in range(3):
    m.encode(wav)
    for code, ms, fl, by, dims in d.get_profile():
        acc[code] = acc.get(code, 0.0) + ms / 3
d.set_profile(False)
names = {1: "gemm", 7: "ws", 8: "slab", 12: "conv0", 13: "act", 14: "ln", 15: "attn", 18: "argmin", 19: "reduce"}
print(f"{os.path.basename(os.path.dirname(sys.argv[1]))}/{os.path.basename(sys.argv[1])}: encode median {ts[len(ts) // 2]:.3f} ms min {ts[0]:.3f}  " +
      " ".join(f"{names.get(k, k)}={v:.2f}" for k, v in sorted(acc.items(), key=lambda kv: -kv[1])))
