"""Developer tool: do a tensor-bound conv and an FP32-bound Activation1d overlap when launched on two streams?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops
B, C, T, k = 16, 384, 13792, 11
torch.manual_seed(0)
x = torch.randn(B, T, C, device="cuda").bfloat16()
w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
bias = torch.randn(C, device="cuda")
xa = torch.randn(B, T, C, device="cuda")
al = 0.5 * torch.randn(C, device="cuda"); be = 0.5 * torch.randn(C, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def conv(): ops.conv1d_nlc(x, w, bias, 1, 1, out_f32=False, out_bf16=True)
def act(): ops.activation1d_nlc(xa, al, be, torch.bfloat16, False)
def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
def both():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1):
        for _ in range(2): conv()
    with torch.cuda.stream(s2):
        for _ in range(6): act()
    cur.wait_stream(s1); cur.wait_stream(s2)
def serial():
    for _ in range(2): conv()
    for _ in range(6): act()
print(f"conv {timed(conv):.4f} ms  act {timed(act):.4f} ms  serial(2 conv + 6 act) {timed(serial):.4f} ms  two streams {timed(both):.4f} ms")
