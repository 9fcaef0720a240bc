#!/bin/bash
# Round-2 pass A: GPU tests after the hygiene batch, TS-mode tcgen05 probe, K1 same-box timing, sanitizer logs.
OUT=gpurun_out/r02a
mkdir -p "${OUT}"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "${OUT}/smi.csv" 2>&1
nproc > "${OUT}/nproc.txt"
timeout 120 ./build/probe_umma_ts > "${OUT}/probe_umma_ts.log" 2>&1; echo "probe rc=$?" | tee -a "${OUT}/probe_umma_ts.log"
cat "${OUT}/probe_umma_ts.log" | tail -45
timeout 900 python -m pytest tests -m gpu -x -q > "${OUT}/pytest.log" 2>&1; echo "pytest rc=$?" | tee -a "${OUT}/pytest.log"
tail -5 "${OUT}/pytest.log"
timeout 600 python tools/bench_act1d.py --quick --k1 --reps 3 --out "${OUT}/act1d_k1_quick.json" > "${OUT}/act1d_k1.log" 2>&1; echo "k1 rc=$?"
tail -14 "${OUT}/act1d_k1.log" | cut -c1-600
# sanitizer: small-shape kernel tests under synccheck and racecheck (mbarrier pipelines of conv_umma / act1d)
SEL='test_conv1d_tcgen05_bf16 or test_convtranspose1d_polyphase or test_activation1d_nlc_multi_equals_single_calls or test_activation1d_nlc_vs_oracle_fp32'
for tool in synccheck racecheck; do
  timeout 900 compute-sanitizer --tool ${tool} --log-file "${OUT}/sanitizer_${tool}.log" python -m pytest tests/test_gpu_kernels.py -x -q -k "${SEL} and not long" > "${OUT}/sanitizer_${tool}_pytest.log" 2>&1
  echo "${tool} rc=$?"; tail -3 "${OUT}/sanitizer_${tool}_pytest.log"; tail -5 "${OUT}/sanitizer_${tool}.log"
done
ls -la "${OUT}"
