"""Developer tool: launch representative layers of the cfg2 workload in isolation (for ncu / event timing).

    python tools/run_layers.py [act|conv|all] [--reps N]
"""
import argparse, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("what", nargs="?", default="all")
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
dev = "cuda"
B = 16
torch.manual_seed(0)

def timeit(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

if args.what in ("act", "all"):
    for C, T in ((384, 13792), (24, 220672)):
        x32 = torch.randn(B, T, C, device=dev)
        x16 = x32.bfloat16()
        al = 0.5 * torch.randn(C, device=dev); be = 0.5 * torch.randn(C, device=dev)
        for name, x in (("f32->bf16", x32), ("bf16->bf16", x16)):
            ms = timeit(lambda: ops.activation1d_nlc(x, al, be, torch.bfloat16, False), args.reps)
            n = B * T * C
            print(f"act nlc C={C} T={T} {name}: {ms:.4f} ms  {n/ms/1e6:.1f} Gelem/s", flush=True)
if args.what in ("conv", "all"):
    for C, T, k, d, res in ((768, 3448, 11, 1, False), (768, 3448, 3, 1, True), (384, 13792, 11, 1, False), (384, 13792, 3, 1, True),
                            (192, 27584, 11, 1, True), (96, 55168, 11, 1, True), (48, 110336, 11, 1, True),
                            (24, 220672, 11, 1, True), (24, 220672, 3, 1, False)):
        x = torch.randn(B, T, C, device=dev).bfloat16()
        w = ops.pack_conv1d_weight(torch.randn(C, C, k, device=dev) / (C * k) ** 0.5, 1)
        bias = torch.randn(C, device=dev)
        r = torch.randn(B, T, C, device=dev) if res else None
        fn = (lambda: ops.conv1d_nlc(x, w, bias, d, 1, residual=r)) if res else (lambda: ops.conv1d_nlc(x, w, bias, d, 1, out_f32=False, out_bf16=True))
        ms = timeit(fn, args.reps)
        fl = 2.0 * C * C * k * B * T
        print(f"conv C={C} T={T} k={k} d={d} res={res}: {ms:.4f} ms  {fl/ms/1e9:.1f} TFLOP/s", flush=True)
