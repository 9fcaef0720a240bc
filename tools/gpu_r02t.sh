#!/bin/bash
# CTA duration spread of the conv kernel on the small-channel layers (trace build): is a dynamic tile scheduler worth it?
OUT=gpurun_out/r02t; mkdir -p $OUT
export BVG_LIB_PATH=build/libbvg_trace.so
for a in "96 55168 11 1" "96 55168 5 1" "96 55168 3 1" "48 110336 11 1" "48 110336 7 1" "24 220672 11 1" "24 220672 7 0" "192 27584 3 1"; do
  echo "== conv C T k res = $a"
  timeout 120 python tools/conv_timeline.py $a 160 2>&1 | grep -E "CTA durations|median period"
done 2>&1 | tee $OUT/cta_spread.log
