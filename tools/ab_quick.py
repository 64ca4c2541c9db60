"""One sample for same-box A/B runs: median device time of the 64 x 10 s bulk step (and optionally the per-kernel table) with a given
build of the library.     python tools/ab_quick.py <path/to/libq3tts_decode.so> [reps] [--kernels]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qwen3_tts_b200 import _capi

_capi.set_library_path(sys.argv[1])
import ctypes
_probe = ctypes.CDLL(sys.argv[1])
for _n in list(_capi.EXPORTS) + list(_capi.DEBUG_EXPORTS):   # an older build under test may lack the newest entry points
    if not hasattr(_probe, _n):
        _capi.EXPORTS.pop(_n, None); _capi.DEBUG_EXPORTS.pop(_n, None)
import torch

from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

reps = int(sys.argv[2]) if len(sys.argv) > 2 and sys.argv[2].isdigit() else 8
cfg = Qwen3TTSTokenizerV2DecoderConfig()
dec = synthetic_model().decoder
codes = make_codes(64, 125, cfg, seed=1).cuda().transpose(1, 2).contiguous()
dec.reserve(64 * 125)
for _ in range(3):
    dec(codes)
ts = []
for _ in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); dec(codes); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
ts.sort()
line = f"{os.path.basename(os.path.dirname(sys.argv[1]))}/{os.path.basename(sys.argv[1])}: step median {ts[len(ts) // 2]:.3f} ms  min {ts[0]:.3f}"
if "--kernels" in sys.argv:
    names = {1: "gemm", 6: "unit", 8: "slab", 7: "ws", 5: "tail"}
    dec.set_profile(True)
    acc = {}
    for _ in range(4):
        dec(codes)
        for code, ms, fl, by, d in dec.get_profile():
            key = (names.get(code, str(code)), d[0], d[1], d[2], d[3])
            acc[key] = acc.get(key, 0.0) + ms / 4
    dec.set_profile(False)
    line += "\n" + "\n".join(f"   {k[0]:5s} rows={k[1]:9d} n={k[2]:5d} cin={k[3]:5d} k={k[4]}  {v:7.3f} ms" for k, v in sorted(acc.items(), key=lambda kv: -kv[1])[:14])
print(line, flush=True)
