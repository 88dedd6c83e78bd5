#!/bin/bash
# usage: gpurun --timeout 900 -- 'bash tools/gpu_edge_round.sh TAG'
TAG=${1:-r01ax}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_edge.py -m gpu -q > $O/${TAG}_pytest_edge.log 2>&1; echo "pytest exit=$?" | tee -a $O/${TAG}_pytest_edge.log
tail -30 $O/${TAG}_pytest_edge.log
timeout 300 python tools/gpu_edge_bench.py > $O/${TAG}_edge_bench.jsonl 2> $O/${TAG}_edge_bench.err; echo "edge bench exit=$?"; cat $O/${TAG}_edge_bench.jsonl; tail -5 $O/${TAG}_edge_bench.err
