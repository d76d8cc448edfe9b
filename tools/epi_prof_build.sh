#!/bin/bash
# Builds tools/_prof/${OUT:-libjm_prof.so}: the library with gemm.cu compiled with -DJM_EPI_PROF (cycle accounting of the
# row-register epilogue, printed by block 0's first epilogue warp).  Needs the normal build's objects.
set -e
cd "$(dirname "$0")/.."
B=jumbo_mae_tpu_b200/build
mkdir -p tools/_prof
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC \
  --expt-relaxed-constexpr ${PROF_DEFS--DJM_EPI_PROF} $EXTRA_DEFS -Iinclude -c jumbo_mae_tpu_b200/csrc/gemm.cu -o tools/_prof/${OUT:-libjm_prof.so}.gemm.o
OBJS=$(ls $B/*.o | grep -v "/gemm.o")
/usr/local/cuda/bin/nvcc -shared -o tools/_prof/${OUT:-libjm_prof.so} tools/_prof/${OUT:-libjm_prof.so}.gemm.o $OBJS \
  -gencode arch=compute_100a,code=sm_100a -lcudart
