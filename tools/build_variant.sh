#!/bin/bash
# Developer tool: build a variant of libbvg.so with extra -D flags for ONE source file, for A/B runs on the GPU box:
#   tools/build_variant.sh act1d "-DACT_TMA_MIN_CTAS=3 -DACT_TMA_PREFETCH=0" build/libbvg_act3.so
#   BVG_LIB_PATH=build/libbvg_act3.so python tools/layer_ab.py
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
SRC="$1"; DEFS="$2"; OUT="$3"
OBJ="${ROOT}/build/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"${NVCC}" -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden --expt-relaxed-constexpr \
  -Xptxas -v ${DEFS} -c "${ROOT}/index-tts_b200/csrc/${SRC}.cu" -o "${OBJ}/${SRC}_var.o" > "${OBJ}/${SRC}_var.log" 2>&1
objs=()
for f in api act1d act1d_umma conv_umma conv_simt pack cond forward; do  # keep in step with csrc/build.sh
  if [ "$f" == "$SRC" ]; then objs+=("${OBJ}/${f}_var.o"); else objs+=("${OBJ}/${f}.o"); fi
done
"${NVCC}" -shared -o "${ROOT}/${OUT}" "${objs[@]}" -lcudart_static -ldl -lrt -lpthread
grep -h "registers\|spill" "${OBJ}/${SRC}_var.log" | sort | uniq -c | sort -rn | head -5
echo "built ${OUT}"
