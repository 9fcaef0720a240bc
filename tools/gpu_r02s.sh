#!/bin/bash
# Co-residency A/B: conv kernel capped at 96 registers (build/libbvg_conv96.so) so that a conv CTA and an Activation1d CTA
# fit one SM together (registers 384*96 + 288*96 <= 64K, shared memory 120 KB + 98 KB), AMP blocks on three streams.
export TAG=r02s
V=build/libbvg_conv96.so
bash tools/ab_bench.sh "BVG_NOP=1" "BVG_LIB_PATH=$V" "BVG_LIB_PATH=$V BVG_CONV_SMEM_KB=120" \
  "BVG_LIB_PATH=$V BVG_CONV_SMEM_KB=120 BVG_ACT_GRID=1" "BVG_LIB_PATH=$V BVG_CONV_SMEM_KB=120 BVG_ACT_GRID=2" \
  "BVG_LIB_PATH=$V BVG_CONV_SMEM_KB=120 BVG_ACT_MULTI=0" "BVG_LIB_PATH=$V BVG_CONV_SMEM_KB=104 BVG_ACT_GRID=1 BVG_ACT_MULTI=0" 2>&1 | tee gpurun_out/r02s/ab.log
