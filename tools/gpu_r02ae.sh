#!/bin/bash
# Outputs per thread of the public [B,C,T] Activation1d kernel: 16 (shipped) vs 24 vs 32
OUT=gpurun_out/r02ae; mkdir -p $OUT
for v in "" build/libbvg_ncl24.so build/libbvg_ncl32.so; do
  echo "== lib ${v:-default (16)}"
  BVG_LIB_PATH=$v timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "ncl or golden" 2>&1 | tail -1
  BVG_LIB_PATH=$v timeout 600 python tools/bench_act1d.py --reps 5 --out $OUT/a.json > $OUT/a.log 2>&1; echo "rc=$?"
  python - $OUT/a.json <<'PY'
import json,sys
j=json.load(open(sys.argv[1]))
for r in j["rows"]:
    if r["C"] in (1536, 96) and r["T"] >= 65536: print(r["dtype"], r["C"], r["T"], "%.3f ms" % r["ms"], "frac %.3f" % r["frac_of_hbm_peak"])
PY
done 2>&1 | tee $OUT/ncl_pt.log
