"""Developer tool: event-timed small-channel convs (stages 3-5 of cfg2) for A/B experiments with BVG_CONV_* knobs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops
B = 16
torch.manual_seed(0)
def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
out = []
for C, T in ((96, 55168), (48, 110336), (24, 220672)):
    for k in (3, 11):
        for res in (False, True):
            x = torch.randn(B, T, C, device="cuda").bfloat16()
            w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
            bias = torch.randn(C, device="cuda")
            r = torch.randn(B, T, C, device="cuda") if res else None
            fn = (lambda: ops.conv1d_nlc(x, w, bias, 1, 1, residual=r)) if res else (lambda: ops.conv1d_nlc(x, w, bias, 1, 1, out_f32=False, out_bf16=True))
            ms = timeit(fn)
            byts = B * T * C * (2 + (4 + 4 if res else 2))
            out.append(f"C{C}k{k}r{int(res)}={ms*1e3:.0f}us({byts/ms/1e6:.0f}GB/s)")
print(" ".join(out), flush=True)
