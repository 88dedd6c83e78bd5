#!/bin/bash
# usage: gpurun --timeout 900 -- 'bash tools/gpu_edge_ncu.sh TAG'   (ncu --set full of the edge kernels, F = 32)
TAG=${1:-r01bc}
O=gpurun_out
mkdir -p $O
CMD="python tools/gpu_edge_bench.py 4194304 32"
$CMD > $O/${TAG}_edge_plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/${TAG}_edge_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'standardize_kernel|moments_partial_vec_kernel|normal_fill_kernel' -s 12 -c 6 -f -o $O/${TAG}_edge $CMD > $O/${TAG}_edge_ncu.log 2>&1
echo "ncu exit=$?"
ncu -i $O/${TAG}_edge.ncu-rep --page raw --csv > $O/${TAG}_edge_ncu_raw.csv 2>/dev/null
python - <<PY
import csv
rows = list(csv.reader(open("$O/${TAG}_edge_ncu_raw.csv")))
hdr = rows[0]
want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active"]
idx = [hdr.index(w) for w in want if w in hdr]
for r in rows[1:]:
    print(" | ".join(r[i][:60] for i in idx))
PY
