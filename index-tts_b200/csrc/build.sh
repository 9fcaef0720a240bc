#!/bin/bash
# Build libbvg.so (sm_100a only) in-tree:  index-tts_b200/libbvg.so
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${HERE}/../libbvg.so"
OBJ="${HERE}/../../build/obj"
mkdir -p "${OBJ}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden --expt-relaxed-constexpr)
pids=()
for f in api act1d act1d_umma conv_umma conv_simt pack cond forward; do
  "${NVCC}" "${FLAGS[@]}" -Xptxas -v -c "${HERE}/${f}.cu" -o "${OBJ}/${f}.o" > "${OBJ}/${f}.log" 2>&1 &
  pids+=($!)
done
rc=0
for p in "${pids[@]}"; do wait "$p" || rc=1; done
if [ $rc -ne 0 ]; then grep -h -E "error|Error" -A3 "${OBJ}"/*.log || cat "${OBJ}"/*.log; exit 1; fi
"${NVCC}" -shared -o "${OUT}" "${OBJ}"/{api,act1d,act1d_umma,conv_umma,conv_simt,pack,cond,forward}.o -lcudart_static -ldl -lrt -lpthread
echo "built ${OUT}"
