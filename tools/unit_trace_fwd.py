"""Timeline of the LAST fused-unit launch of one bulk forward (CTA 0, first 16 tiles): Q3TTS_DBG=32 is set here.
    python tools/unit_trace_fwd.py [B] [T]        (C = 96, dilation 9, raw-input form unless Q3TTS_NO_XIN=1)"""
import ctypes as C
import os
import sys

os.environ["Q3TTS_DBG"] = "32"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
# The GemmParams::dbg experiment bits and the Q3TTS_* switches exist only in the -DQ3TTS_DEBUG build:
#     python -m qwen3_tts_b200.build --debug      ->  qwen3_tts_b200/lib/libq3tts_decode_dbg.so
from qwen3_tts_b200 import _capi, build as _build
_capi.set_library_path(_build.LIB_PATH_DEBUG)
import torch

from qwen3_tts_b200 import _capi, synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 125
cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model()
dec = model.decoder
codes = make_codes(B, T, cfg, seed=1).cuda().transpose(1, 2).contiguous()
for _ in range(2):
    dec(codes)
torch.cuda.synchronize()
lib = _capi.load_library()
tr = (C.c_longlong * 256)()
lib.q3tts_unit_trace(tr)
names = ["M1 start", "-", "M1 issued", "a2_full", "acc2_empty", "M2 issued", "P1 start", "P1 done", "P2 start", "P2 done", "S start", "S done", "S wait", "P1 ld", "P1 math+st", "P1 st wait"]
t0 = tr[0]
print("tile " + " ".join(f"{n[:10]:>10}" for n in names))
for it in range(16):
    print(f"{it:4d} " + " ".join(f"{(tr[it*16+e]-t0) if tr[it*16+e] else 0:10d}" for e in range(16)))
