#!/bin/bash
# A/B of the CTA-pair (multicast) variant of the layer-wise kernels; each run under its own timeout (a hang must not eat the box)
BFK_LW_CLUSTER=2 timeout 120 python -m pytest tests/test_gpu_layerwise.py -q --tb=short -x 2>&1 | tail -4
echo "cluster=2"; BFK_LW_CLUSTER=2 B=606208 STEPS=5 timeout 120 python tools/gpu_lw_prof.py | tail -1
echo "cluster=1"; BFK_LW_CLUSTER=1 B=606208 STEPS=5 timeout 120 python tools/gpu_lw_prof.py | tail -1
