#!/bin/bash
# training step time of every build/libbfk_<name>.so (bench.py's cfg5_fm_train through that library):
#   bash tools/gpu_train_variant.sh [steps]
for so in build/libbfk_*.so; do
  echo "== $so"
  for i in 1 2; do
    BFK_LIB=$so timeout 200 python -c "
import os, sys, runpy
import bayesflow_b200._lib as L
L.LIB_PATH = os.path.abspath(os.environ['BFK_LIB'])
sys.argv = ['bench.py', '--workload', 'cfg5_fm_train', '--steps', '${1:-30}', '--warmup', '5']
runpy.run_path('bench.py', run_name='__main__')" 2>&1 | grep -o '"ms_per_step": [0-9.]*' | head -1
  done
done
