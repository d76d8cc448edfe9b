#!/bin/bash
# Builds experiment variants of the library into exp/lib_<name>.so:  tools/build_variants.sh name1="-DX" name2="-DY -DZ"
set -e
cd "$(dirname "$0")/.."
mkdir -p exp
for spec in "$@"; do
  name="${spec%%=*}"; defs="${spec#*=}"
  touch jumbo_mae_tpu_b200/csrc/*.cu
  JM_NVCC_DEFS="$defs" python -m jumbo_mae_tpu_b200.build > /dev/null 2> exp/build_$name.log
  cp jumbo_mae_tpu_b200/libjumbo_mae_b200.so exp/lib_$name.so
  echo "built $name ($defs)"
done
touch jumbo_mae_tpu_b200/csrc/*.cu
python -m jumbo_mae_tpu_b200.build > /dev/null 2>&1
