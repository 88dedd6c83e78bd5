#!/bin/bash
# One gpurun call: the whole GPU suite -> smoke -> every bench line.  usage: gpurun --timeout 1800 -- 'bash tools/gpu_full_round.sh TAG'
TAG=${1:-r01y}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit=$?" | tee -a $O/${TAG}_pytest.log
tail -8 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke exit=$?"; tail -2 $O/${TAG}_smoke.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit=$?"; cut -c1-2500 $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
timeout 600 python bench.py --workload cfg4_wide_fm_euler100 --steps 2 --warmup 3 --no-train --no-cpu-baseline > $O/${TAG}_bench_cfg4.json 2>> $O/${TAG}_bench.err; echo "bench cfg4 exit=$?"; cut -c1-700 $O/${TAG}_bench_cfg4.json
timeout 600 python bench.py --workload cfg5_fm_train --steps 10 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_train.json 2>> $O/${TAG}_bench.err; echo "bench train exit=$?"; cut -c1-700 $O/${TAG}_bench_train.json
timeout 600 python bench.py --workload cfg5_fm_train_wide --steps 5 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_train_wide.json 2>> $O/${TAG}_bench.err; echo "bench wide exit=$?"; cut -c1-700 $O/${TAG}_bench_train_wide.json
tail -5 $O/${TAG}_bench.err
