"""Developer tool: event times of CTA 0's FIRST tile in the tcgen05 conv kernel (bvg_debug_conv_trace), for layers where
every CTA owns a single tile (one utterance at stages 0-1).   python tools/conv_first_tile.py B C T k"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops, _lib
B, C, T, k = (int(a) for a in sys.argv[1:5])
torch.manual_seed(0)
x = torch.randn(B, T, C, device="cuda").bfloat16()
w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
bias = torch.randn(C, device="cuda")
fn = lambda: ops.conv1d_nlc(x, w, bias, 1, 1, out_f32=False, out_bf16=True)
fn(); fn(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); fn(); e1.record(); torch.cuda.synchronize()
N = 8
buf = torch.zeros(16 * N, dtype=torch.int64, device="cuda")
lib = _lib.load()
lib.bvg_debug_conv_trace(buf.data_ptr(), N)
fn(); torch.cuda.synchronize()
lib.bvg_debug_conv_trace(None, 0)
t = buf.cpu().view(16, N)
names = {0: "prod a_empty ok", 1: "mma acc_empty ok", 2: "mma a_full ok", 3: "mma committed", 4: "epi acc_full ok", 5: "epi tmem read", 6: "epi tile done"}
base = int(t[t > 0].min())
print(f"B={B} C={C} T={T} k={k}: kernel {1e3 * e0.elapsed_time(e1):.1f} us (warm L2); k-steps per tile {k * C // 16}")
for s, n in names.items():
    vals = [int(v) - base for v in t[s].tolist() if v > 0]
    print(f"  {n:18s} {vals}")
