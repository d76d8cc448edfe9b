#!/bin/bash
# builds the library with each experiment switch and times the attention kernels (run on the GPU box)
for defs in "-DJM_EXP_NO_STATS" "-DJM_EXP_NO_SOFTMAX" "-DJM_EXP_NO_SOFTMAX -DJM_EXP_NO_STATS" "-DJM_EXP_NO_SOFTMAX -DJM_EXP_NO_STATS -DJM_EXP_NO_DRAIN"; do
  JM_NVCC_DEFS="$defs" python -m jumbo_mae_tpu_b200.build > /dev/null 2>&1
  echo "== $defs"
  timeout 120 python tools/attn_check.py both 2>&1 | tail -3 | head -2
done
