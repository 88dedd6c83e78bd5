#!/bin/bash
# usage: gpurun --timeout 900 -- 'bash tools/gpu_ot_round.sh TAG'
TAG=${1:-r01bd}
O=gpurun_out
mkdir -p $O
timeout 800 python -m pytest tests/test_gpu_ot.py -m gpu -q > $O/${TAG}_pytest_ot.log 2>&1; echo "pytest exit=$?" | tee -a $O/${TAG}_pytest_ot.log
tail -60 $O/${TAG}_pytest_ot.log
