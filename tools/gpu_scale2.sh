#!/bin/bash
# 2-GPU check of both bench arms (sampling shards by dataset; training all-reduces gradients over NCCL).
TAG=${1:-r01s2}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_n2.json 2> $O/${TAG}_n2.err; echo "n2 exit=$?"; cut -c1-600 $O/${TAG}_bench_n2.json; tail -3 $O/${TAG}_n2.err
timeout 600 $TR bench.py --gpus 2 --workload cfg5_fm_train --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_train_n2.json 2>> $O/${TAG}_n2.err; echo "train n2 exit=$?"; cut -c1-600 $O/${TAG}_train_n2.json
timeout 600 $TR bench.py --gpus 2 --workload cfg5_fm_train_wide --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_train_wide_n2.json 2>> $O/${TAG}_n2.err; echo "train wide n2 exit=$?"; cut -c1-600 $O/${TAG}_train_wide_n2.json
timeout 600 $TR bench.py --gpus 2 --workload cfg4_wide_fm_euler100 --steps 2 --warmup 3 --no-cpu-baseline --no-train > $O/${TAG}_cfg4_n2.json 2>> $O/${TAG}_n2.err; echo "cfg4 n2 exit=$?"; cut -c1-300 $O/${TAG}_cfg4_n2.json
timeout 300 $TR bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > $O/${TAG}_ref_n2.json 2>> $O/${TAG}_n2.err; echo "ref n2 exit=$?"; cut -c1-400 $O/${TAG}_ref_n2.json
tail -5 $O/${TAG}_n2.err
