#!/bin/bash
# How many warps does the direct-load (no TMA) Activation1d need?  CTAs per SM capped by padding dynamic shared memory.
OUT=gpurun_out/r02r; mkdir -p $OUT
for kb in 0 70 100 200; do
  echo "== BVG_ACT_DIRECT=1 pad ${kb} KB"
  BVG_ACT_DIRECT=1 BVG_ACT_DIRECT_SMEM_KB=$kb timeout 300 python tools/act_umma_check.py --time --skip-check 2>/dev/null | grep '"C"' | cut -c1-110
done 2>&1 | tee $OUT/act_direct_occupancy.log
