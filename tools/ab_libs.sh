#!/bin/bash
# Same-box A/B of two builds of the library, alternating A B A B A B in fresh processes:  bash tools/ab_libs.sh libA.so libB.so [extra args of ab_quick.py]
A=$1; B=$2; shift 2
for i in 1 2 3; do
  timeout 300 python tools/ab_quick.py $A 8 "$@" 2>&1 | tail -16
  timeout 300 python tools/ab_quick.py $B 8 "$@" 2>&1 | tail -16
done
