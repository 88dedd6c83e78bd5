#!/bin/bash
# Training round: parity tests (training + density share the forward), both training bench lines, A/B against the unfused forward.
TAG=${1:-r01am}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_density.py tests/test_gpu_gemm.py -q --tb=line > $O/${TAG}_pytest.log 2>&1; echo "pytest exit=$?"; tail -12 $O/${TAG}_pytest.log | cut -c1-400
python bench.py --workload cfg5_fm_train --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_train.json 2> $O/${TAG}_bench.err; echo "bench exit=$?"; cut -c1-200 $O/${TAG}_bench_train.json
python bench.py --workload cfg5_fm_train_wide --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_train_wide.json 2>> $O/${TAG}_bench.err; echo "bench wide exit=$?"; cut -c1-200 $O/${TAG}_bench_train_wide.json
BFK_TRAIN_FUSED=0 python bench.py --workload cfg5_fm_train --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | cut -c1-200
tail -3 $O/${TAG}_bench.err
