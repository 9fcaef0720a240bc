#!/bin/bash
# Cluster multicast of the weight stages for the pair-mode layers (BVG_CONV_CLUSTER=3): parity + per-layer timing
OUT=gpurun_out/r02ab; mkdir -p $OUT
for c in 0 3; do
  BVG_CONV_CLUSTER=$c timeout 300 python tools/conv_cluster_check.py 2>&1 | grep -E "C=96|ALL|FAIL" 
done | tee $OUT/cluster_pair.log
for c in 0 3; do
  echo "== layer_ab BVG_CONV_CLUSTER=$c"; BVG_CONV_CLUSTER=$c timeout 300 python tools/layer_ab.py 4 2>/dev/null | grep -E "conv1d  st 3|conv1d  st 4|conv1d  st 5|conv1d  st 2|total|conv"
done | tee $OUT/layer_ab.log
