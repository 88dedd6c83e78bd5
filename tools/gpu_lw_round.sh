#!/bin/bash
TAG=${1:-r01w}
O=gpurun_out
mkdir -p $O
python -m pytest tests/test_gpu_layerwise.py -x -q > $O/${TAG}_pytest_lw.log 2>&1; echo "pytest lw exit=$?"; tail -8 $O/${TAG}_pytest_lw.log
python -m pytest tests/test_gpu_train.py tests/test_gpu_gemm.py -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest train exit=$?"; tail -3 $O/${TAG}_pytest.log
python bench.py --workload cfg5_fm_train --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_train.json 2> $O/${TAG}_bench.err; echo "bench exit=$?"; cut -c1-200 $O/${TAG}_bench_train.json
timeout 600 python bench.py --workload cfg4_wide_fm_euler100 --steps 2 --warmup 1 --no-train > $O/${TAG}_bench_cfg4.json 2>> $O/${TAG}_bench.err; echo "bench cfg4 exit=$?"; cut -c1-1800 $O/${TAG}_bench_cfg4.json; tail -3 $O/${TAG}_bench.err
