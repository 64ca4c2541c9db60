"""Summarise an .ncu-rep with source counters: opcode mix, stall reasons, instruction groups by execution count.
    python tools/ncu_sass_summary.py report.ncu-rep [min_total_M]"""
import collections
import csv
import io
import re
import subprocess
import sys

rep = sys.argv[1]
min_total = float(sys.argv[2]) * 1e6 if len(sys.argv) > 2 else 3e6
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, v = rows[0], rows[2]
for key in ("gpu__time_duration.sum", "launch__registers_per_thread", "launch__block_size", "smsp__inst_executed.sum",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__t_bytes.sum", "sm__cycles_elapsed.max"):
    for i, a in enumerate(h):
        if a.endswith(key):
            print(f"{a} = {v[i]} {rows[1][i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h, data = rows[1], rows[2:]
iS, iI, iN = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
tot = sum(int(r[iI]) for r in data)
totS = sum(int(r[iN]) for r in data)
print("total warp instructions", tot, "samples", totS)
ops, smp = collections.Counter(), collections.Counter()
def opof(s):
    m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", s)
    return m.group(2) if m else "?"
for r in data:
    op = opof(r[iS]).split(".")[0]
    ops[op] += int(r[iI]); smp[op] += int(r[iN])
for op, c in ops.most_common(24):
    print(f"  {op:10} {c:12d} {100*c/tot:5.1f}%  samples {100*smp[op]/max(totS,1):5.1f}%")
for name in h:
    if name.startswith("stall_") and "Not Issued" not in name:
        i = h.index(name)
        s = sum(int(r[i]) for r in data)
        if s > totS * 0.02:
            print(f"  {name} {100*s/totS:.1f}%")
prev, start, cnt, sm, oo = None, 0, 0, 0, []
def flush():
    if prev is not None and prev * cnt > min_total:
        print(f"--- idx {start:5d} n={cnt:4d} exec={prev:10d} total={prev*cnt/1e6:8.1f}M samples={sm:6d} :: " + " ".join(oo[:12]))
for k, r in enumerate(data):
    c = int(r[iI])
    if prev is None or abs(c - prev) > 0.02 * max(prev, 1):
        flush(); prev, start, cnt, sm, oo = c, k, 0, 0, []
    cnt += 1; sm += int(r[iN]); oo.append(opof(r[iS]))
flush()
