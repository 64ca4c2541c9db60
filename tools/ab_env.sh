#!/bin/bash
# Same-box A/B of one library under two environments: tools/ab_env.sh "ENV_A=1" "ENV_B=1"   (empty string = no variable)
set -u
for i in 1 2; do
  for E in "$1" "$2"; do
    echo "== env: ${E:-<none>}"
    env $E timeout 300 python tools/profile_layers.py 64 125 2>&1 | python -c "
import sys, collections
tot = collections.OrderedDict(); head = None
for ln in sys.stdin:
    f = ln.split()
    if ln.startswith('B='): head = ln.strip(); continue
    if len(f) == 10 and f[0].isdigit():
        k = (f[1], f[3], f[4], f[5]); tot[k] = tot.get(k, 0.0) + float(f[6])
print(head)
print(' '.join(f'{k[0]}{k[1]}x{k[2]}k{k[3]}={v:.2f}' for k, v in tot.items()))
"
  done
done
