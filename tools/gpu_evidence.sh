#!/bin/bash
# one gpurun call: GPU suite, smoke, both bench arms, ncu launch list + one full capture of the dominant kernel
# usage (on the GPU box, from the repo root): bash tools/gpu_evidence.sh r02e
tag=${1:-r02x}
out=gpurun_out
mkdir -p $out
python -m pytest tests -q -m gpu > $out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/${tag}_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/${tag}_smoke.log
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_ref.json 2> $out/${tag}_ref.err; echo "ref rc=$?"
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
tail -c 1500 $out/${tag}_bench.json
# the other BASELINE.json configurations (parity-test cases, not bench lines): one line each for BASELINE.md's table
for wl in cfg1_twomoons_fm_euler100 cfg2_sir_fm_rk45x100 cfg3_lv_dm_euler100 cfg4_wide_fm_euler100; do
  python bench.py --workload $wl --steps 3 --warmup 3 --no-train --no-cpu-baseline > $out/${tag}_bench_${wl}.json 2>/dev/null
  python - "$out/${tag}_bench_${wl}.json" <<'PY'
import json, sys
b = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(b["config"]["workload"], "%.3f M draws/s" % (b["value"] / 1e6), "e2e %.3f M" % (b["e2e"]["value"] / 1e6), "frac %.3f" % b["roofline"]["frac"], b["clocks"]["reasons"])
PY
done
short="python bench.py --steps 2 --warmup 1 --no-train --no-cpu-baseline"
$short > $out/${tag}_plain.log 2>&1 || exit 0
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $short > $out/${tag}_ncu_launch.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sampler_fold -s 1 -c 1 -f -o $out/${tag}_fold $short > $out/${tag}_ncu_full.log 2>&1
echo "full capture rc=$?"
ls -la $out | tail -12
