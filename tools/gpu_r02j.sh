#!/bin/bash
OUT=gpurun_out/r02j
mkdir -p "${OUT}"
for d in 0 1 2 4 8 3 6 7 15; do
  echo "== dbg=$d"
  BVG_ACT_DBG=$d timeout 200 python tools/act_umma_check.py --time --skip-check 2>/dev/null | grep '"C": 384\|"C": 24' | cut -c1-120
done > "${OUT}/dbg.log" 2>&1
cat "${OUT}/dbg.log"
