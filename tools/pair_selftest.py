"""CTA-pair (cta_group::2) building-block check and MMA rate.  usage: pair_selftest.py"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qwen3_tts_b200 import _capi
lib = _capi.load_library()
out = (C.c_float * 4)()
for n, k, ts in ((96, 96 // 32 * 64 if False else 64, 0), (96, 192, 0), (192, 192, 0), (256, 256, 0), (64, 64, 1), (128, 128, 1)):
    rc = lib.q3tts_selftest_pair(0, n, k, 0, ts, out)
    print(f"n={n} k={k} ts={ts} rc={rc} maxdiff D {out[0]:.3e} D2 {out[1]:.3e}", flush=True)
for n in (96, 192, 256):
    rc = lib.q3tts_selftest_pair(0, n, 64, 2000, 0, out)
    print(f"rate n={n}: rc={rc} {out[2]:.1f} clk per M=256 K=16 MMA", flush=True)
