"""tcgen05.mma cycles per instruction vs N / operand form / swizzle / A row shift.  python tools/mma_bench.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qwen3_tts_b200 import _capi

lib = _capi.load_library()
print("   N span shift form nacc commit wait  clk/MMA   MAC/clk")
for n in (96, 192, 256):
    for form, mode in ((0, 0), (1, 0), (0, 1), (0, 2), (0, 4), (0, 7), (0, -1)):
        v = C.c_float(0)
        rc = lib.q3tts_bench_mma(0, n, 128, 0, form, 1, mode, 1, C.byref(v))
        print(f"N={n:4d} form={form} mode={mode:2d}  clk/MMA {v.value:8.1f}  MAC/clk {128*n*16/max(v.value,1e-9):7.0f} rc={rc}", flush=True)
