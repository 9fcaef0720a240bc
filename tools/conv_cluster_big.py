"""Developer tool: cluster-multicast A/B on the big weight-streaming layers of cfg2 (B = 16).  BVG_CONV_CLUSTER=0|2"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops
B = 16
torch.manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for C, T, k, d, res in ((768, 3448, 3, 1, True), (768, 3448, 11, 1, False), (384, 13792, 3, 1, True), (384, 13792, 7, 1, True),
                        (384, 13792, 11, 1, False), (192, 27584, 3, 1, True), (192, 27584, 7, 1, True), (192, 27584, 11, 1, True),
                        (192, 27584, 11, 5, False)):
    x = torch.randn(B, T, C, device="cuda").bfloat16()
    w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
    bias = torch.randn(C, device="cuda")
    r = torch.randn(B, T, C, device="cuda") if res else None
    fn = (lambda: ops.conv1d_nlc(x, w, bias, d, 1, residual=r)) if res else (lambda: ops.conv1d_nlc(x, w, bias, d, 1, out_f32=False, out_bf16=True))
    y0 = fn(); y1 = fn(); torch.cuda.synchronize()
    ms = []
    for _ in range(7):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    print(f"cluster={os.environ.get('BVG_CONV_CLUSTER', '0')} C={C} T={T} k={k} d={d} res={int(res)}: {1e3 * ms[len(ms) // 2]:.1f} us  "
          f"checksum {float(y0.float().abs().sum()):.6e} repeat-equal {bool(torch.equal(y0, y1))}", flush=True)
