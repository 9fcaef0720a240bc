#!/bin/bash
# Round-2 final pass: full check (tests, bench + layer table, isolated layers, ncu launch list and captures), cfg3 microbench,
# DRAM-traffic launch list, reference arm.
bash tools/gpu_check.sh r02ad
OUT=gpurun_out/r02ad
timeout 900 python tools/bench_act1d.py --reps 5 --out "${OUT}/act1d_microbench_cfg3.json" > "${OUT}/act1d_microbench.log" 2>&1; echo "cfg3 microbench rc=$?"
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-v1"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2400 --csv \
  --log-file "${OUT}/launches_traffic_all.csv" $CMD > "${OUT}/ncu_traffic.log" 2>&1
echo "ncu traffic rc=$?"
cp profiles/ncu_traffic.json "${OUT}/ncu_traffic.json"
python tools/launch_traffic.py "${OUT}/launches_traffic_all.csv" "${OUT}/ncu_traffic.json" > "${OUT}/launch_traffic.txt"; tail -3 "${OUT}/launch_traffic.txt"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > "${OUT}/bench_reference.json" 2> "${OUT}/bench_reference.err"; echo "bench ref rc=$?"
grep -E "passed|failed" "${OUT}/pytest.log"
