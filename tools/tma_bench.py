"""TMA box-shape microbenchmark table.  python tools/tma_bench.py"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qwen3_tts_b200 import _capi

lib = _capi.load_library()
print("C  box(cols x rows)  dir    warps  GB/s")
for Cc, bc, br in [(96, 32, 32), (96, 32, 64), (96, 32, 128), (96, 96, 32), (96, 96, 16), (192, 32, 32), (192, 64, 32), (192, 64, 16), (192, 64, 64), (192, 192, 32),
                   (1024, 64, 32), (1024, 64, 128)]:
    for load in (0, 1):
        for warps in (8, 16):
            g = C.c_float(0)
            rc = lib.q3tts_bench_tma(0, Cc, bc, br, load, warps, 4000, C.byref(g))
            print(f"{Cc:5d} {bc:4d} x {br:4d}   {'load ' if load else 'store'} {warps:3d}   {g.value:8.1f}  rc={rc}", flush=True)
