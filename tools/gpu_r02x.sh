#!/bin/bash
# fp32 public op: vectorised unaligned-row path.  Parity (odd T, strided input, cfg3 full size), then the cfg3 quick shapes.
OUT=gpurun_out/r02x; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "ncl or golden" > $OUT/pytest_ncl.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_ncl.log
timeout 600 python tools/bench_act1d.py --quick --reps 5 --out $OUT/act1d_quick.json > $OUT/act1d_quick.log 2>&1; echo "rc=$?"
python - $OUT/act1d_quick.json <<'PY'
import json,sys
j=json.load(open(sys.argv[1]))
for r in j["rows"]: print(r["dtype"], r["C"], r["T"], "%.3f ms" % r["ms"], "%.0f GB/s" % r["achieved_GBs"], "frac %.3f" % r["frac_of_hbm_peak"])
PY
