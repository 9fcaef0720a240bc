#!/bin/bash
# fp32 only: 20 / 28 outputs per thread (conflict-free 80- / 112-byte lane strides) against the shipped 24
OUT=gpurun_out/r02af; mkdir -p $OUT
for v in "" build/libbvg_ncl20.so build/libbvg_ncl28.so; do
  echo "== lib ${v:-default (24)}"
  BVG_LIB_PATH=$v timeout 300 python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
from index_tts_b200 import ops
import numpy as np
from oracle import bigvgan_oracle as O
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
rng = np.random.default_rng(0)
x = (rng.standard_normal((2, 6, 50000)) * 1.5).astype(np.float32)
al = (0.5 * rng.standard_normal(6)).astype(np.float32); be = (0.5 * rng.standard_normal(6)).astype(np.float32)
y = ops.activation1d(torch.from_numpy(x).cuda(), torch.from_numpy(al).cuda(), torch.from_numpy(be).cuda(), True).cpu().numpy()
print('max err vs oracle', float(np.abs(y - O.activation1d(x.astype(np.float64), al, be)).max()))
for T in (2 ** 20, 2 ** 21, 2 ** 20 + 37):
    xc = torch.randn(1, 1536, T, device='cuda'); a = 0.5 * torch.randn(1536, device='cuda'); b = 0.5 * torch.randn(1536, device='cuda')
    for _ in range(2): ops.activation1d(xc, a, b, False)
    ms = []
    for _ in range(5):
        flush.fill_(1); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.activation1d(xc, a, b, False); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    m = sorted(ms)[2]
    print('f32', T, '%.3f ms' % m, 'frac %.3f' % (2 * 4 * xc.numel() / m / 1e6 / 6540.5))
    del xc
PY
done 2>&1 | tee $OUT/ncl_pt_f32.log
