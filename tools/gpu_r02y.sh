#!/bin/bash
OUT=gpurun_out/r02y; mkdir -p $OUT
for dt in f32 bf16; do
  C1="python tools/run_one.py ncl 1536 262144 $dt"
  timeout 120 $C1 > $OUT/plain_ncl_$dt.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:act1d_ncl -s 2 -c 1 -f -o $OUT/prof_ncl_c1536_$dt $C1 > $OUT/ncu_ncl_$dt.log 2>&1
  echo "ncu ncl $dt rc=$?"
done
