#!/bin/bash
# compute-sanitizer (memcheck / racecheck / synccheck) over CI-sized shapes of the hand-rolled mbarrier / TMEM kernels.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_sanitize.sh TAG'
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
python tools/gpu_sanitize.py all > $O/${TAG}_sanitize_plain.log 2>&1; echo "plain exit=$?"; tail -12 $O/${TAG}_sanitize_plain.log
for tool in memcheck synccheck racecheck; do
  for part in fused lw train; do
    timeout 420 compute-sanitizer --tool $tool --print-limit 20 python tools/gpu_sanitize.py $part > $O/${TAG}_sanitize_${tool}_${part}.log 2>&1
    echo "$tool $part exit=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize driver done|Error|hazard" $O/${TAG}_sanitize_${tool}_${part}.log | head -6
  done
done
