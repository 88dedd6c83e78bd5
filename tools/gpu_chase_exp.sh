#!/bin/bash
# run every build/libbfk_<name>.so variant through tools/gpu_chase_prof.py (cfg2 shape, Euler-20 scaled to 100 evals)
for so in build/libbfk_*.so; do
  n=$(basename $so .so); n=${n#libbfk_}
  echo "=== $n"
  BFK_LIB=$so timeout 120 python tools/gpu_chase_prof.py ${1:-20} ${2:-1000} 2>&1 | grep -v Warning | tail -${3:-2}
done
