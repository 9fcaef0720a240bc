#!/bin/bash
# One GPU-box pass: parity tests, bench line + per-launch table, isolated layer timings, ncu launch list and
# one full ncu capture of the two dominant kernels.  Usage (from the repo root): gpurun -- bash tools/gpu_check.sh [tag]
TAG="${1:-run}"
OUT=gpurun_out/${TAG}
mkdir -p "${OUT}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "${OUT}/smi.csv" 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > "${OUT}/pytest.log" 2>&1; echo "pytest rc=$?" | tee -a "${OUT}/pytest.log"
tail -5 "${OUT}/pytest.log"
timeout 600 python bench.py --steps 5 --warmup 3 --layer-table "${OUT}/layers.json" > "${OUT}/bench.json" 2> "${OUT}/bench.err"; echo "bench rc=$?"
cat "${OUT}/bench.json" | cut -c1-1500
timeout 300 python tools/run_layers.py all --reps 5 > "${OUT}/layers_isolated.log" 2>&1; echo "layers rc=$?"
cat "${OUT}/layers_isolated.log"
if [ "${SKIP_NCU:-0}" != "1" ]; then
  CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
  timeout 300 $CMD > "${OUT}/plain.log" 2>&1 &&
  timeout 700 ncu --metrics gpu__time_duration.sum --clock-control none -c 1800 --csv --log-file "${OUT}/launches_all.csv" $CMD > "${OUT}/ncu_launches.log" 2>&1
  echo "ncu launches rc=$?"
  python tools/launch_share.py "${OUT}/launches_all.csv" "${OUT}/launches.csv" > "${OUT}/launches_summary.txt"; cat "${OUT}/launches_summary.txt"
  C1="python tools/run_one.py conv 384 13792 11 1 0"
  timeout 120 $C1 > "${OUT}/plain_conv.log" 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 -f -o "${OUT}/prof_conv_c384k11" $C1 > "${OUT}/ncu_conv.log" 2>&1
  echo "ncu conv rc=$?"
  C2="python tools/run_one.py act 384 13792 f32"
  timeout 120 $C2 > "${OUT}/plain_act.log" 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:act1d -s 2 -c 1 -f -o "${OUT}/prof_act_c384" $C2 > "${OUT}/ncu_act.log" 2>&1
  echo "ncu act rc=$?"
fi
ls -la "${OUT}"
if [ "${SKIP_NCU:-0}" != "1" ]; then
  C3="python tools/run_one.py act 24 220672 bf16"
  timeout 120 $C3 > "${OUT}/plain_act24.log" 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:act1d -s 2 -c 1 -f -o "${OUT}/prof_act_c24_bf16" $C3 > "${OUT}/ncu_act24.log" 2>&1
  echo "ncu act24 rc=$?"
  C4="python tools/run_one.py conv 24 220672 11 1 1"
  timeout 120 $C4 > "${OUT}/plain_conv24.log" 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_umma -s 2 -c 1 -f -o "${OUT}/prof_conv_c24k11res" $C4 > "${OUT}/ncu_conv24.log" 2>&1
  echo "ncu conv24 rc=$?"
fi
ls -la "${OUT}"
