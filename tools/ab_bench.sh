#!/bin/bash
# On the GPU box: runs bench.py once per variant library in exp/ (args: variant names), alternating.
cd "$(dirname "$0")/.."
cp jumbo_mae_tpu_b200/libjumbo_mae_b200.so exp/lib_keep.so
for v in "$@"; do
  cp exp/lib_$v.so jumbo_mae_tpu_b200/libjumbo_mae_b200.so
  echo "== $v: $(timeout 250 python bench.py --no-cpu-baseline 2>&1 | tail -1 | cut -c1-100)"
done
cp exp/lib_keep.so jumbo_mae_tpu_b200/libjumbo_mae_b200.so
