#!/bin/bash
# Round-2 pass U: full check with the new tests and bench legs, DRAM-traffic launch list, reference arm.
SKIP_NCU=1 bash tools/gpu_check.sh r02u
OUT=gpurun_out/r02u
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-v1"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2400 --csv \
  --log-file "${OUT}/launches_traffic_all.csv" $CMD > "${OUT}/ncu_traffic.log" 2>&1
echo "ncu traffic rc=$?"
python tools/launch_traffic.py "${OUT}/launches_traffic_all.csv" "${OUT}/ncu_traffic.json" > "${OUT}/launch_traffic.txt"; cat "${OUT}/launch_traffic.txt"
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > "${OUT}/bench_reference.json" 2> "${OUT}/bench_reference.err"; echo "bench ref rc=$?"
cut -c1-300 "${OUT}/bench_reference.json"
grep -E "passed|failed" "${OUT}/pytest.log"
