#!/bin/bash
# Round-2 pass B: split-operand fp32 mode tests + TS probe timing + fp32-mode latency
OUT=gpurun_out/r02b
mkdir -p "${OUT}"
timeout 120 ./build/probe_umma_ts > "${OUT}/probe_umma_ts.log" 2>&1; echo "probe rc=$?" | tee -a "${OUT}/probe_umma_ts.log"
tail -12 "${OUT}/probe_umma_ts.log"
timeout 1200 python -m pytest tests -m gpu -x -q > "${OUT}/pytest.log" 2>&1; echo "pytest rc=$?" | tee -a "${OUT}/pytest.log"
tail -25 "${OUT}/pytest.log"
timeout 300 python - > "${OUT}/fp32_latency.log" 2>&1 <<'PY'
import sys, time, torch
sys.path.insert(0, '.')
import bench
m = bench.make_state_dict_torch().cuda()
mel = torch.randn(1, 80, 431, device='cuda')
res = {}
for prec in ('fp32', 'fp32_simt', 'bf16'):
    m.set_precision(prec)
    with torch.no_grad():
        for _ in range(3): w = m(mel)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5): w = m(mel)
        e1.record(); torch.cuda.synchronize()
    res[prec] = w.clone()
    print(prec, 'cfg1 ms', e0.elapsed_time(e1) / 5)
d = (res['fp32'] - res['fp32_simt']).abs().max().item()
print('max |split - simt| =', d, ' max|wav| =', res['fp32_simt'].abs().max().item())
PY
cat "${OUT}/fp32_latency.log" | grep -v Warn | tail -6
