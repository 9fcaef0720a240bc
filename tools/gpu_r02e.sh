#!/bin/bash
OUT=gpurun_out/r02e
mkdir -p "${OUT}"
timeout 120 ./build/probe_umma_ts > "${OUT}/probe_umma_ts.log" 2>&1; echo "probe rc=$?"
grep -i "accumulator\|issue rate\|SUMMARY" "${OUT}/probe_umma_ts.log"
timeout 600 python tools/act_umma_check.py --time > "${OUT}/umma.log" 2>&1; echo "umma rc=$?" | tee -a "${OUT}/umma.log"
grep -v " ok$" "${OUT}/umma.log" | tail -30
for io in f32 bf16; do
  C2="python tools/run_one.py act 384 13792 ${io}"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:act1d_umma -s 2 -c 1 -f -o "${OUT}/prof_act_umma_c384_${io}" $C2 > "${OUT}/ncu_act_${io}.log" 2>&1
  echo "ncu ${io} rc=$?"
done
ls -la "${OUT}"
