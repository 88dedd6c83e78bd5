#!/bin/bash
# usage: gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_shard_round.sh TAG'
TAG=${1:-r01bg}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519"
timeout 300 $TR tools/gpu_shard_check.py > $O/${TAG}_shard_check.log 2>&1; echo "shard check exit=$?"; grep -E "PASS|FAIL|Error" $O/${TAG}_shard_check.log | head; tail -3 $O/${TAG}_shard_check.log
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_n2.json 2> $O/${TAG}_n2.err; echo "n2 exit=$?"; cut -c1-300 $O/${TAG}_bench_n2.json; tail -3 $O/${TAG}_n2.err
