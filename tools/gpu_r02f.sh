#!/bin/bash
OUT=gpurun_out/r02i
mkdir -p "${OUT}"
timeout 600 python tools/act_umma_check.py --time > "${OUT}/umma.log" 2>&1; echo "umma rc=$?" | tee -a "${OUT}/umma.log"
grep -v " ok$" "${OUT}/umma.log" | tail -30
C2="python tools/run_one.py act 384 13792 f32"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:act1d_umma -s 2 -c 1 -f -o "${OUT}/prof_act_umma_c384_f32" $C2 > "${OUT}/ncu_act_f32.log" 2>&1
echo "ncu rc=$?"
