#!/bin/bash
# one gpurun call after a change to the TRAINING kernels only: GPU suite, smoke, both bench arms, ncu launch lists of the two
# training workloads (the sampler captures of tools/gpu_evidence.sh stand).  usage: bash tools/gpu_evidence_train.sh r02an
tag=${1:-r02x}
out=gpurun_out
mkdir -p $out
python -m pytest tests -q -m gpu > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${tag}_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_ref.json 2> $out/${tag}_ref.err; echo "ref rc=$?"
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 1500 $out/${tag}_bench.json
for wl in cfg5_fm_train cfg5_fm_train_wide; do
  python bench.py --workload $wl --steps 20 --warmup 5 > $out/${tag}_bench_${wl}.json 2>/dev/null; echo "$wl rc=$?"
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $out/${tag}_${wl}_launches.csv \
    python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > $out/${tag}_${wl}_ncu.log 2>&1
  echo "$wl launch list rc=$?"
done
python tools/gpu_train_small.py 8192 300 1 > $out/${tag}_train8192.log 2>&1
ls -la $out | tail -12
