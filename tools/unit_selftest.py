"""Fused residual-unit kernel vs the SIMT cross-check, one subprocess per case (a trapped launch poisons the context).
    python tools/unit_selftest.py"""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CASES = [(1000, 96, 1, 0), (1000, 96, 9, 0), (5000, 96, 3, 0), (700, 192, 1, 0), (700, 192, 9, 0), (3000, 192, 3, 1), (255, 96, 9, 1),
         (1, 96, 1, 0), (129, 192, 3, 0), (257, 96, 3, 0), (40000 * 148 // 100, 96, 1, 0), (40000, 96, 9, 0), (40000, 192, 9, 0)]

if len(sys.argv) > 1:
    from qwen3_tts_b200 import _capi
    lib = _capi.load_library()
    rows, ch, dil, bf = (int(v) for v in sys.argv[1:5])
    out = (C.c_float * 4)()
    rc = lib.q3tts_selftest_unit(0, rows, ch, dil, bf, out)
    print(f"rows={rows} C={ch} dil={dil} bf16={bf}: rc={rc} y_diff={out[0]:.4e} y_ref={out[1]:.3f} s_diff={out[2]:.4e} s_ref={out[3]:.3f}")
    sys.exit(0 if rc == 0 else 1)

for case in CASES:
    r = subprocess.run([sys.executable, __file__, *map(str, case)], capture_output=True, text=True, timeout=300)
    print(r.stdout.strip() or f"{case}: no output", flush=True)
    if r.returncode != 0:
        print("  stderr:", r.stderr.strip()[-400:], flush=True)
