"""Developer tool: board power and SM clock while looping one kernel type for ~3 s each."""
import os, subprocess, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops
B, C, T, k = 16, 384, 13792, 11
x = torch.randn(B, T, C, device="cuda").bfloat16()
w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
bias = torch.randn(C, device="cuda")
xa = torch.randn(B, T, C, device="cuda")
al = 0.5 * torch.randn(C, device="cuda"); be = 0.5 * torch.randn(C, device="cuda")
x24 = torch.randn(B, 220672, 24, device="cuda").bfloat16()
w24 = ops.pack_conv1d_weight(torch.randn(24, 24, 11, device="cuda") / 16, 1)
b24 = torch.randn(24, device="cuda"); r24 = torch.randn(B, 220672, 24, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def conv(): ops.conv1d_nlc(x, w, bias, 1, 1, out_f32=False, out_bf16=True)
def act(): ops.activation1d_nlc(xa, al, be, torch.bfloat16, False)
def conv24(): ops.conv1d_nlc(x24, w24, b24, 1, 1, residual=r24)
def both():
    with torch.cuda.stream(s1): conv()
    with torch.cuda.stream(s2):
        for _ in range(3): act()
def sample(tag, fn, secs=3.0):
    rows = []
    p = subprocess.Popen(["nvidia-smi", "--query-gpu=power.draw,clocks.sm", "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
    th = threading.Thread(target=lambda: [rows.append(l) for l in p.stdout], daemon=True); th.start()
    t0 = time.time(); n = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < secs:
        for _ in range(20): fn()
        n += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    p.terminate()
    vals = [tuple(float(v) for v in r.split(",")) for r in rows[5:] if "," in r]
    pw = sorted(v[0] for v in vals); ck = sorted(v[1] for v in vals)
    print(f"{tag:10s} {e0.elapsed_time(e1)/n:.4f} ms/iter  power median {pw[len(pw)//2]:.0f} W max {pw[-1]:.0f}  clock median {ck[len(ck)//2]:.0f} MHz")
sample("idle", lambda: None, 1.0)
sample("conv384", conv); sample("act384", act); sample("conv24", conv24); sample("conv+3act", both)
