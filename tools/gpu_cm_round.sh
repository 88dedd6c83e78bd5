#!/bin/bash
# usage: gpurun --timeout 900 -- 'bash tools/gpu_cm_round.sh TAG'
TAG=${1:-r01ba}
O=gpurun_out
mkdir -p $O
timeout 800 python -m pytest tests/test_gpu_consistency.py tests/test_gpu_edge.py -m gpu -q -x > $O/${TAG}_pytest_cm.log 2>&1; echo "pytest exit=$?" | tee -a $O/${TAG}_pytest_cm.log
tail -40 $O/${TAG}_pytest_cm.log
