#!/bin/bash
# channels-last Activation1d: 3 CTAs/SM variant and slab-size knobs against the default, stage shapes of cfg2
OUT=gpurun_out/r02aj; mkdir -p $OUT
run() { echo "== $*"; env "$@" timeout 200 python tools/act_umma_check.py --time --skip-check 2>/dev/null | grep '"C"' | python -c "
import sys,json
for l in sys.stdin:
    r=json.loads(l); print(r['C'], r['io'], round(r['us'],1))" | paste - - - - - - ; }
run BVG_NOP=1
run BVG_LIB_PATH=build/libbvg_act3.so BVG_ACT_SMEM_KB=72
run BVG_LIB_PATH=build/libbvg_act3.so BVG_ACT_SMEM_KB=60
run BVG_ACT_RB=1
run BVG_ACT_RB=2
run BVG_ACT_RB=4
run BVG_ACT_SMEM_KB=64
