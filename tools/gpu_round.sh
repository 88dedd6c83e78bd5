#!/bin/bash
# One gpurun call: parity tests -> bench -> ncu launch list -> ncu full capture of the sampler kernel.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_round.sh TAG'
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest exit=$?" | tee -a $O/${TAG}_pytest.log
tail -5 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke exit=$?"; tail -2 $O/${TAG}_smoke.log
python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit=$?"; cat $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
for wl in cfg2_sir_fm_rk45x100 cfg3_lv_dm_euler100 cfg1_twomoons_fm_euler100; do
  python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > $O/${TAG}_bench_$wl.json 2>> $O/${TAG}_bench.err; cat $O/${TAG}_bench_$wl.json | cut -c1-400
done
SHORT="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
$SHORT > $O/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${TAG}_launches.csv $SHORT > $O/${TAG}_ncu1.log 2>&1
echo "ncu launches exit=$?"
$SHORT > $O/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:sampler_ -s 1 -c 1 -f -o $O/${TAG}_sampler $SHORT > $O/${TAG}_ncu2.log 2>&1
echo "ncu full exit=$?"
ls -la $O
