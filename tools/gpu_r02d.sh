#!/bin/bash
OUT=gpurun_out/r02d
mkdir -p "${OUT}"
timeout 600 python tools/act_umma_check.py --time > "${OUT}/umma.log" 2>&1; echo "umma rc=$?" | tee -a "${OUT}/umma.log"
cat "${OUT}/umma.log" | tail -70
BVG_ACT_UMMA=0 timeout 300 python tools/act_umma_check.py --time --skip-check > "${OUT}/simt.log" 2>&1; echo "simt rc=$?"
cat "${OUT}/simt.log" | tail -14
