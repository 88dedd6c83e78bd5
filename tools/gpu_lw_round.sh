#!/bin/bash
# Layer-wise (cfg4) round: parity tests, throughput probe, warm launch list, one ncu --set full capture of the block kernel.
TAG=${1:-r01ar}
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_layerwise.py tests/test_gpu_stochastic.py -q --tb=short > $O/${TAG}_pytest_lw.log 2>&1; echo "pytest lw exit=$?"; tail -4 $O/${TAG}_pytest_lw.log | cut -c1-300
B=606208 STEPS=5 python tools/gpu_lw_prof.py | tail -1
B=75776 STEPS=2 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --cache-control none --clock-control none -c 300 --csv --log-file $O/${TAG}_lw_launches.csv python tools/gpu_lw_prof.py > $O/${TAG}_ncu_list.log 2>&1; echo "ncu list exit=$?"
B=75776 STEPS=2 ncu --set full --clock-control none --import-source on -k regex:lw_block_gemm -s 14 -c 1 -f -o $O/${TAG}_lw_block python tools/gpu_lw_prof.py > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full exit=$?"; tail -2 $O/${TAG}_ncu_full.log
