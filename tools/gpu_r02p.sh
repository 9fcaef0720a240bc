#!/bin/bash
# Round-2 pass P: full check (tests, bench + layer table, launch list, ncu captures) + whole-step A/B of the opt-in
# tensor-core Activation1d kernel + the reference arm.
bash tools/gpu_check.sh r02p
OUT=gpurun_out/r02p
BVG_ACT_UMMA=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-v1 > "${OUT}/bench_act_umma1.json" 2> "${OUT}/bench_act_umma1.err"; echo "bench umma rc=$?"
cut -c1-400 "${OUT}/bench_act_umma1.json"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > "${OUT}/bench_reference.json" 2> "${OUT}/bench_reference.err"; echo "bench ref rc=$?"
cat "${OUT}/bench_reference.json" | cut -c1-1200
