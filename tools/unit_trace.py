"""Timeline of the fused unit kernel (CTA 0, first 16 tiles).  Q3TTS_DBG=32 python tools/unit_trace.py rows C dil"""
import ctypes as C
import os
import sys

os.environ["Q3TTS_DBG"] = str(32 | int(os.environ.get("UT_EXTRA", "0")))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
# The GemmParams::dbg experiment bits and the Q3TTS_* switches exist only in the -DQ3TTS_DEBUG build:
#     python -m qwen3_tts_b200.build --debug      ->  qwen3_tts_b200/lib/libq3tts_decode_dbg.so
from qwen3_tts_b200 import _capi, build as _build
_capi.set_library_path(_build.LIB_PATH_DEBUG)

lib = _capi.load_library()
rows, ch, dil = (int(v) for v in sys.argv[1:4])
out = (C.c_float * 4)()
rc = lib.q3tts_selftest_unit(0, rows, ch, dil, 0, out)
tr = (C.c_longlong * 256)()
lib.q3tts_unit_trace(tr)
names = ["M1 start", "slab0 ready", "M1 issued", "a2_full seen", "acc2_empty seen", "M2 issued", "P1 start(acc1_full)", "P1 done", "P2 start(acc2_full)", "P2 done"]
t0 = tr[0]
print(f"rows={rows} C={ch} dil={dil} rc={rc} diff={out[0]:.3e}")
print("tile " + " ".join(f"{n[:12]:>12}" for n in names))
for it in range(int(os.environ.get('UT_ROWS', '16'))):
    print(f"{it:4d} " + " ".join(f"{tr[it*16+e]-t0:12d}" for e in range(10)))
