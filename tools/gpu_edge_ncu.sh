#!/bin/bash
# usage: gpurun --timeout 900 -- 'bash tools/gpu_edge_ncu.sh TAG'   (ncu --set full, one launch of each edge kernel, F = 32)
TAG=${1:-r01bf}
O=gpurun_out
mkdir -p $O
CMD="python tools/gpu_edge_bench.py 4194304 32"
$CMD > $O/${TAG}_edge_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/${TAG}_edge_plain.log; exit 1; }
for K in standardize_kernel moments_partial_vec_kernel normal_fill_kernel; do
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 1 -f -o $O/${TAG}_$K $CMD > $O/${TAG}_${K}_ncu.log 2>&1
  echo "ncu $K exit=$?"
  ncu -i $O/${TAG}_$K.ncu-rep --page raw --csv > $O/${TAG}_${K}_raw.csv 2>/dev/null
  echo "== $K"; python tools/ncu_brief.py $O/${TAG}_$K.ncu-rep 2>&1 | grep -E "gpu__time|dram__bytes|issue_active.avg.pct|registers|warps_active|pipe_xu|pipe_alu|pipe_fma\\.|lts__throughput"
done
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench exit=$?"; python - <<PY
import json
d = json.loads(open("$O/${TAG}_bench.json").read().strip().splitlines()[-1])
print(json.dumps({k: d.get(k) for k in ("value", "e2e", "sampler_api", "edge_kernels")}, indent=1)[:3000])
PY
tail -3 $O/${TAG}_bench.err
