"""Developer tool: one isolated launch for ncu.  python tools/run_one.py conv C T k d res | act C T in_dtype | ncl C T dtype"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops
B = 16
torch.manual_seed(0)
kind = sys.argv[1]
if kind == "conv":
    C, T, k, d, res = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
    x = torch.randn(B, T, C, device="cuda").bfloat16()
    w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
    bias = torch.randn(C, device="cuda")
    r = torch.randn(B, T, C, device="cuda") if res else None
    for _ in range(3):
        if res: ops.conv1d_nlc(x, w, bias, d, 1, residual=r)
        else: ops.conv1d_nlc(x, w, bias, d, 1, out_f32=False, out_bf16=True)
elif kind == "ncl":  # the public [B, C, T] op, fast-sine mode
    C, T = int(sys.argv[2]), int(sys.argv[3])
    x = torch.randn(1, C, T, device="cuda")
    if sys.argv[4] == "bf16": x = x.bfloat16()
    al = 0.5 * torch.randn(C, device="cuda"); be = 0.5 * torch.randn(C, device="cuda")
    for _ in range(3):
        ops.activation1d(x, al, be, False)
else:
    C, T = int(sys.argv[2]), int(sys.argv[3])
    x = torch.randn(B, T, C, device="cuda")
    if sys.argv[4] == "bf16": x = x.bfloat16()
    al = 0.5 * torch.randn(C, device="cuda"); be = 0.5 * torch.randn(C, device="cuda")
    for _ in range(3):
        ops.activation1d_nlc(x, al, be, torch.bfloat16, False)
torch.cuda.synchronize()
print("done")
