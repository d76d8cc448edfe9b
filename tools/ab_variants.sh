#!/bin/bash
# On the GPU box: runs the GEMM micro-bench once per variant library in exp/ (args: variant names; FILTER env = grep pattern)
cd "$(dirname "$0")/.."
cp jumbo_mae_tpu_b200/libjumbo_mae_b200.so /tmp/lib_keep.so
for v in "$@"; do
  cp exp/lib_$v.so jumbo_mae_tpu_b200/libjumbo_mae_b200.so
  echo "== $v"
  ONLY=${ONLY:-gemm} timeout 120 python tools/bench_kernels.py 2>&1 | grep -v "timed out" | grep -E "${FILTER:-total}"
done
cp /tmp/lib_keep.so jumbo_mae_tpu_b200/libjumbo_mae_b200.so
