#!/bin/bash
OUT=gpurun_out/r02ak; mkdir -p $OUT
run() { echo "== $*"; env "$@" timeout 200 python tools/act_umma_check.py --time --skip-check 2>/dev/null | grep '"C"' | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['C'], r['io'], round(r['us'],1))" | paste - - - - - - ; }
run BVG_NOP=1
run BVG_LIB_PATH=build/libbvg_actpf.so
run BVG_LIB_PATH=build/libbvg_actpf.so BVG_ACT_RB=1
run BVG_LIB_PATH=build/libbvg_actpf.so BVG_ACT_RB=4
