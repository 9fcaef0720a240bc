#!/bin/bash
# GPU pass: parity tests, then A/B of the dual-pipeline conv (BVG_CONV_GROUPS) and concurrent AMP blocks (BVG_STREAMS=1 = serial)
TAG="${1:-ab}"; OUT=gpurun_out/${TAG}; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest.log
tail -4 $OUT/pytest.log
for g in 1 2; do
  echo "== isolated convs BVG_CONV_GROUPS=$g"
  BVG_CONV_GROUPS=$g timeout 300 python tools/run_layers.py conv --reps 5 2>&1 | tee $OUT/layers_g$g.log
done
TAG=$TAG bash tools/ab_bench.sh "BVG_CONV_GROUPS=1 BVG_STREAMS=1" "BVG_CONV_GROUPS=1 BVG_STREAMS=0" "BVG_CONV_GROUPS=2 BVG_STREAMS=1" "BVG_CONV_GROUPS=2 BVG_STREAMS=0" 2>&1 | tee $OUT/ab.log
