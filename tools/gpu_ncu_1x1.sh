#!/bin/bash
# ncu --set full of the first 1x1 conv of the C = 384 residual units (general tcgen05 GEMM kernel with residual + raw + Snake'd outputs).
set -u
mkdir -p gpurun_out
timeout 300 python tools/ncu_target.py 64 125 1 > gpurun_out/target.log 2>&1 || { echo target failed; tail -5 gpurun_out/target.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_fwd.csv python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_fwd.log 2>&1
SKIP=$(python - <<'PY'
import csv
rows = [r for r in csv.reader(open('gpurun_out/launches_fwd.csv')) if len(r) > 10]
h = rows[0]; iK = h.index('Kernel Name'); iV = h.index('Metric Value')
n = -1; pick = None; lst = []
for r in rows[1:]:
    if 'conv_gemm_tc_kernel' in r[iK]:
        n += 1
        ms = float(r[iV].replace(',', '')) / 1e6
        lst.append((n, ms))
        if pick is None and 0.7 < ms < 1.5: pick = n
open('gpurun_out/gemm_launches.txt', 'w').write('\n'.join(f'{a} {b:.4f}' for a, b in lst))
print(pick if pick is not None else 0)
PY
)
echo "skip=$SKIP"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_gemm_tc_kernel -s $SKIP -c 1 -o gpurun_out/prof_1x1_384 -f \
    python tools/ncu_target.py 64 125 1 > gpurun_out/ncu_1x1.log 2>&1; echo "ncu rc=$?"
python tools/ncu_sass_summary.py gpurun_out/prof_1x1_384.ncu-rep 8 > gpurun_out/ncu_1x1_384.txt 2>&1
head -70 gpurun_out/ncu_1x1_384.txt
