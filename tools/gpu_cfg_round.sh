#!/bin/bash
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_cfg_round.sh TAG'  (the bench lines of configs 1-3 with the current kernels)
TAG=${1:-r01bm}
O=gpurun_out
mkdir -p $O
for wl in cfg1_twomoons_fm_euler100 cfg2_sir_fm_rk45x100 cfg3_lv_dm_euler100; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --no-train > $O/${TAG}_bench_$wl.json 2>> $O/${TAG}_bench.err; echo "$wl exit=$?"
  python - <<PY
import json
d = json.loads(open("$O/${TAG}_bench_$wl.json").read().strip().splitlines()[-1])
print(json.dumps({"workload": d["config"]["workload"], "value": d["value"], "e2e": d["e2e"]["value"], "frac": d["roofline"]["frac"], "cpu": d.get("cpu_baseline")})[:600])
PY
done
tail -3 $O/${TAG}_bench.err
