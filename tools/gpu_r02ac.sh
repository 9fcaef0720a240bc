#!/bin/bash
# Dynamic tile scheduler of the conv kernel: parity (whole suite), per-layer A/B, step A/B.
OUT=gpurun_out/r02ac; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_forward.py -m gpu -x -q > $OUT/pytest_fwd.log 2>&1; echo "pytest fwd rc=$?"; tail -3 $OUT/pytest_fwd.log
for c in 0 1; do
  echo "== layer_ab BVG_CONV_DYNAMIC=$c"; BVG_CONV_DYNAMIC=$c timeout 300 python tools/layer_ab.py 4 2>/dev/null | grep -E "conv|total"
done | tee $OUT/layer_ab.log
export TAG=r02ac
bash tools/ab_bench.sh "BVG_CONV_DYNAMIC=0" "BVG_CONV_DYNAMIC=1" "BVG_CONV_DYNAMIC=0" "BVG_CONV_DYNAMIC=1" | tee $OUT/ab.log
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest.log
