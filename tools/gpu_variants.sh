#!/bin/bash
# every build/libbfk_<name>.so through tools/gpu_variant_check.py: usage  bash tools/gpu_variants.sh [steps] [datasets] [act]
for so in build/libbfk_*.so; do
  n=$(basename $so .so); n=${n#libbfk_}
  echo "=== $n"
  BFK_LIB=$so timeout 180 python tools/gpu_variant_check.py ${1:-20} ${2:-1000} ${3:-mish} 2>&1 | grep -v Warning | tail -1
done
