"""BASELINE.json configs[2]: the Activation1d microbench (public [B, C, T] op, isolated, HBM-bound by its algorithmic
bytes 2*s*C*T).  C = 1536 -> 24, T up to 2^21 samples, fp32 and bf16 I/O, plus one odd length.

    python tools/bench_act1d.py [--out profiles/rNN_act1d_microbench.json] [--quick]

Prints one line per case and writes a JSON table: achieved algorithmic GB/s against the measured HBM copy peak
(MEASURED_PEAKS.json).  CUDA events on the launching stream, inputs larger than L2 or an L2 flush in between.
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
ap.add_argument("--quick", action="store_true")
ap.add_argument("--reps", type=int, default=5)
args = ap.parse_args()

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    peak = 6551.7
dev = torch.device("cuda")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
Cs = (1536, 768, 384, 192, 96, 48, 24)
Ts = (2 ** 14, 2 ** 16, 2 ** 18, 2 ** 20, 2 ** 21, 2 ** 20 + 37)
if args.quick:
    Cs, Ts = (1536, 96, 24), (2 ** 16, 2 ** 20 + 37)
rows = []
for dt in (torch.float32, torch.bfloat16):
    for C in Cs:
        al = 0.5 * torch.randn(C, device=dev)
        be = 0.5 * torch.randn(C, device=dev)
        for T in Ts:
            x = torch.randn(1, C, T, device=dev, dtype=torch.float32).to(dt)
            for _ in range(2):
                y = ops.activation1d(x, al, be, False)
            torch.cuda.synchronize()
            ms = 0.0
            for _ in range(args.reps):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                y = ops.activation1d(x, al, be, False)
                e1.record()
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1)
            ms /= args.reps
            nbytes = 2.0 * x.element_size() * C * T
            gbs = nbytes / (ms * 1e-3) / 1e9
            row = {"dtype": "f32" if dt == torch.float32 else "bf16", "C": C, "T": T, "ms": ms, "algorithmic_bytes": nbytes,
                   "achieved_GBs": gbs, "frac_of_hbm_peak": gbs / peak, "finite": bool(torch.isfinite(y[..., ::4097]).all())}
            rows.append(row)
            print(json.dumps(row), flush=True)
            del x, y
if args.out:
    json.dump({"workload": "cfg3: Activation1d microbench, public [1, C, T] op, fast-sin mode", "hbm_peak_GBs": peak, "l2": "256 MiB flush write "
               "between timed launches", "rows": rows}, open(args.out, "w"), indent=1)
