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
ap.add_argument("--k1", action="store_true", help="also time the reference's fused extension K1 (baseline/_ref/k1, built for "
                "sm_100 by baseline/install_ref.sh) and the reference's unfused torch Activation1d on the same tensors")
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
k1, k1_why, ref_act = None, None, None
if args.k1:
    from baseline import ref_loader
    k1, k1_why = ref_loader.load_k1()
    print(json.dumps({"k1": "loaded" if k1 is not None else f"unavailable: {k1_why}"}), flush=True)
    try:
        _, ref_activations, RefActivation1d = ref_loader.import_reference()
        ref_act = (ref_activations, RefActivation1d)
    except Exception as e:  # pragma: no cover
        print(json.dumps({"reference_torch_path": f"unavailable: {e}"}), flush=True)
    from index_tts_b200.alias_free_activation.torch.filter import kaiser_sinc_filter1d
    ftr = kaiser_sinc_filter1d(0.25, 0.3, 12).to(dev)  # the [1, 1, 12] buffer both resamplers register


def time_fn(fn):
    for _ in range(2):
        y = fn()
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(args.reps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        y = fn()
        e1.record()
        torch.cuda.synchronize()
        ms += e0.elapsed_time(e1)
    return ms / args.reps, y


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
            if args.k1 and C * T < 2 ** 31 - 2 ** 20:  # K1 indexes with int32 (anti_alias_activation_cuda.cu:68-74)
                if k1 is not None:
                    try:
                        # reference call shape: activation1d.py:26-28 (filters are the module buffers; 2x gain on the
                        # up filter is applied inside the kernel, .cu:107)
                        ms1, y1 = time_fn(lambda: k1.forward(x, ftr.to(dt), ftr.to(dt), al.to(dt), be.to(dt)))
                        row["k1_ms"] = ms1
                        row["k1_GBs"] = nbytes / (ms1 * 1e-3) / 1e9
                        row["speedup_vs_k1"] = ms1 / ms
                        row["max_abs_vs_k1"] = float((y1.float() - y.float()).abs().max())
                        del y1
                    except Exception as e:
                        row["k1_error"] = f"{type(e).__name__}: {e}"[:200]
                if ref_act is not None and C * T <= 2 ** 27:  # unfused torch path (act.py:25-30): 3 launches + pads
                    acts, RA = ref_act
                    mod = RA(activation=acts.SnakeBeta(C, alpha_logscale=True)).to(dev)
                    with torch.no_grad():
                        mod.act.alpha.copy_(al); mod.act.beta.copy_(be)
                        mod = mod.to(dt)
                        ms2, y2 = time_fn(lambda: mod(x))
                    row["torch_unfused_ms"] = ms2
                    row["speedup_vs_torch_unfused"] = ms2 / ms
                    row["max_abs_vs_torch_unfused"] = float((y2.float() - y.float()).abs().max())
                    del y2, mod
            rows.append(row)
            print(json.dumps(row), flush=True)
            del x, y
if args.out:
    json.dump({"workload": "cfg3: Activation1d microbench, public [1, C, T] op, fast-sin mode", "hbm_peak_GBs": peak, "l2": "256 MiB flush write "
               "between timed launches", "rows": rows}, open(args.out, "w"), indent=1)
