#!/bin/bash
# ncu --set full capture of the sampler kernel on the short bench (after the same command exited 0 without ncu)
TAG=${1:-prof}
O=gpurun_out
mkdir -p $O
SHORT="python bench.py --steps 1 --warmup 1 --no-cpu-baseline ${2:-}"
$SHORT > $O/${TAG}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sampler_ -s 1 -c 1 -f -o $O/${TAG}_sampler $SHORT > $O/${TAG}_ncu.log 2>&1
echo "ncu full exit=$?"; cut -c1-300 $O/${TAG}_plain.log
