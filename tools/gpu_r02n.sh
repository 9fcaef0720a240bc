#!/bin/bash
# Round-2 pass N (after the machine was replaced): the no-alloc entry point alone and after the suite, the probes whose
# logs were lost, Activation1d SIMT vs tensor-core timings per stage width, K1 same-box timing, fp32-mode latency.
OUT=gpurun_out/r02n
mkdir -p "${OUT}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "${OUT}/smi.csv" 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s > "${OUT}/pytest_noalloc_alone.log" 2>&1; echo "noalloc alone rc=$?" | tee -a "${OUT}/pytest_noalloc_alone.log"
grep -E "noalloc|passed|failed|Error" "${OUT}/pytest_noalloc_alone.log" | tail -12
timeout 1500 python -m pytest tests -m gpu -x -q -s > "${OUT}/pytest.log" 2>&1; echo "pytest rc=$?" | tee -a "${OUT}/pytest.log"
grep -E "noalloc|passed|failed|FAILED|Error" "${OUT}/pytest.log" | tail -14
timeout 120 ./build/probe_umma_ts > "${OUT}/probe_umma_ts.log" 2>&1; echo "probe ts rc=$?" | tee -a "${OUT}/probe_umma_ts.log"
timeout 120 ./build/probe_mma_sync_rate > "${OUT}/probe_mma_sync_rate.log" 2>&1; echo "probe mma rc=$?" | tee -a "${OUT}/probe_mma_sync_rate.log"
tail -6 "${OUT}/probe_mma_sync_rate.log"
for u in 0 1; do
  BVG_ACT_UMMA=$u timeout 300 python tools/act_umma_check.py --time --skip-check 2>/dev/null | grep '"C"' > "${OUT}/act_time_umma${u}.jsonl"
  echo "== BVG_ACT_UMMA=$u"; cut -c1-110 "${OUT}/act_time_umma${u}.jsonl"
done
timeout 300 python tools/act_umma_check.py > "${OUT}/act_umma_check.log" 2>&1; echo "act check rc=$?"; tail -2 "${OUT}/act_umma_check.log"
timeout 600 python tools/bench_act1d.py --quick --k1 --reps 3 --out "${OUT}/act1d_k1_quick.json" > "${OUT}/act1d_k1.log" 2>&1; echo "k1 rc=$?"
tail -14 "${OUT}/act1d_k1.log" | cut -c1-400
ls -la "${OUT}"
