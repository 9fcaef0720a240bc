"""Developer tool: parity + timing of the tcgen05 conv on SMALL-M wide layers (the shapes one utterance produces at stages
0-1), for the cluster-multicast A/B:  BVG_CONV_CLUSTER=0|1 python tools/conv_cluster_check.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from index_tts_b200 import ops
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ok = True
for (B, C, T, k, d, res) in [(1, 768, 1724, 11, 1, False), (1, 768, 1724, 3, 5, True), (1, 768, 1600, 7, 3, True), (1, 384, 6896, 11, 1, False),
                             (1, 384, 6896, 7, 1, True), (2, 768, 700, 11, 5, False), (1, 768, 130, 3, 1, False), (1, 384, 257, 11, 3, True),
                             (16, 96, 55168, 11, 1, True), (16, 96, 55168, 7, 3, False), (3, 96, 1000, 11, 5, True), (16, 96, 55168, 7, 1, True)]:
    g = torch.Generator().manual_seed(C + T + k)
    x = torch.randn(B, T, C, generator=g).cuda().bfloat16()
    w = (torch.randn(C, C, k, generator=g) / (C * k) ** 0.5).cuda()
    bias = torch.randn(C, generator=g).cuda()
    r = torch.randn(B, T, C, generator=g).cuda() if res else None
    wp = ops.pack_conv1d_weight(w, 1)
    fn = lambda: ops.conv1d_nlc(x, wp, bias, d, 1, residual=r)
    y = fn()
    ref = F.conv1d(x.float().transpose(1, 2), w.bfloat16().float(), bias, dilation=d, padding=(k - 1) // 2 * d).transpose(1, 2)
    if res:
        ref = ref + r
    err = float((y - ref).abs().max())
    y2 = fn()
    same = bool(torch.equal(y, y2))
    ms = []
    for _ in range(10):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    good = err < 2e-3 and same
    ok = ok and good
    print(f"cluster={os.environ.get('BVG_CONV_CLUSTER', '0')} B={B} C={C} T={T} k={k} d={d} res={int(res)}: max-abs {err:.2e} repeat-equal {same} "
          f"cold {1e3 * ms[len(ms) // 2]:.1f} us {'OK' if good else 'FAIL'}")
print("ALL OK" if ok else "FAILED")
