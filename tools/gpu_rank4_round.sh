#!/bin/bash
# usage: gpurun --timeout 1200 -- 'bash tools/gpu_rank4_round.sh TAG'
TAG=${1:-r01bi}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_ot.py tests/test_gpu_consistency.py -m gpu -q > $O/${TAG}_pytest_rank4.log 2>&1; echo "pytest exit=$?"; tail -15 $O/${TAG}_pytest_rank4.log | cut -c1-300
timeout 600 python tools/gpu_rank4_bench.py > $O/${TAG}_rank4_bench.jsonl 2> $O/${TAG}_rank4_bench.err; echo "rank4 bench exit=$?"; cat $O/${TAG}_rank4_bench.jsonl; tail -4 $O/${TAG}_rank4_bench.err
