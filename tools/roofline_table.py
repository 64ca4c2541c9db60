"""Per-stage roofline fractions from a tools/profile_layers.py table.
    python tools/roofline_table.py profiles/r1_layers_bulk_64x10s.txt [tflops_peak] [hbm_gbs_peak]
bound time of a launch = max(algorithmic FLOPs / tensor peak, minimum HBM bytes / HBM peak); frac = bound time / measured time."""
import json
import os
import sys

path = sys.argv[1]
peaks = {}
pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.isfile(pk):
    peaks = json.load(open(pk))
tf = float(sys.argv[2]) if len(sys.argv) > 2 else float(peaks.get("bf16_tflops_sustained", peaks.get("tflops_sustained", 1407.7)))
bw = float(sys.argv[3]) if len(sys.argv) > 3 else float(peaks.get("hbm_gbs", 6553.9))
groups = {}
order = []
for line in open(path):
    f = line.split()
    if len(f) < 10 or not f[0].isdigit():
        continue
    op, rows, n, cin, k, ms, tfl, gbs = f[1], int(f[2]), int(f[3]), int(f[4]), int(f[5]), float(f[6]), float(f[8]), float(f[9])
    key = (op, rows, n, cin, k)
    if key not in groups:
        groups[key] = [0, 0.0, 0.0, 0.0]
        order.append(key)
    g = groups[key]
    g[0] += 1; g[1] += ms; g[2] += tfl * ms; g[3] += gbs * ms          # flops (TFLOP/s*ms = GFLOP), bytes (GB/s*ms = MB)
print(f"peaks: tensor {tf:.1f} TFLOP/s (sustained), HBM {bw:.1f} GB/s")
print(f"{'stage':34s} {'n':>2} {'ms':>8} {'TFLOP/s':>8} {'GB/s':>7} {'bound':>6} {'frac':>5}")
tot = tb = 0.0
for key in order:
    cnt, ms, gf, mb = groups[key]
    if ms <= 0:
        continue
    t_tensor = gf / tf          # GFLOP / (TFLOP/s) = ms
    t_hbm = mb / bw             # MB / (GB/s) = ms
    bound = max(t_tensor, t_hbm)
    name = f"{key[0]} rows={key[1]} n={key[2]} cin={key[3]} k={key[4]}"
    tot += ms; tb += bound
    if ms >= 0.05:
        print(f"{name:34s} {cnt:2d} {ms:8.3f} {gf/ms:8.1f} {mb/ms:7.0f} {'tensor' if t_tensor >= t_hbm else 'hbm':>6} {bound/ms:5.2f}")
print(f"{'all stages':34s}    {tot:8.3f} {'':8s} {'':7s} {'':>6} {tb/tot:5.2f}")
