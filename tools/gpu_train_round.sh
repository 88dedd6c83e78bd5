#!/bin/bash
# One gpurun call for the training path: parity tests -> training bench -> ncu launch list of one training step.
TAG=${1:-r01t}
O=gpurun_out
mkdir -p $O
python -m pytest tests/test_gpu_train.py tests/test_gpu_gemm.py -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit=$?"; tail -8 $O/${TAG}_pytest.log
python bench.py --workload cfg5_fm_train --steps 10 --warmup 3 > $O/${TAG}_bench_train.json 2> $O/${TAG}_bench.err; echo "bench exit=$?"; cut -c1-1500 $O/${TAG}_bench_train.json; tail -3 $O/${TAG}_bench.err
python bench.py --workload cfg5_fm_train_wide --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_train_wide.json 2>> $O/${TAG}_bench.err; echo "bench wide exit=$?"; cut -c1-600 $O/${TAG}_bench_train_wide.json
SHORT="python bench.py --workload cfg5_fm_train --steps 1 --warmup 1 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu1.log 2>&1
echo "ncu launches exit=$?"
