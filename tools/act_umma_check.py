"""Developer tool: the tensor-core Activation1d kernel (csrc/act1d_umma.cu) against the NumPy oracle (small shapes) and the
range-reduced SIMT kernel (large shapes), plus CUDA-event timings of the stage shapes of cfg2.

    python tools/act_umma_check.py [--time]          (run with BVG_ACT_UMMA=1; without it the SIMT kernel is timed)
"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from index_tts_b200 import ops
from oracle import bigvgan_oracle as O

ap = argparse.ArgumentParser()
ap.add_argument("--time", action="store_true")
ap.add_argument("--skip-check", action="store_true")
a = ap.parse_args()
dev = torch.device("cuda")
rng = np.random.default_rng(3)


def snr(ref, y):
    ref = ref.astype(np.float64); y = y.astype(np.float64)
    return 10 * np.log10((ref ** 2).sum() / max(((ref - y) ** 2).sum(), 1e-300))


bad = 0
if not a.skip_check:
    for (B, C, T) in [(1, 128, 32), (1, 128, 64), (2, 8, 1), (1, 8, 5), (2, 24, 7), (1, 24, 333), (2, 48, 1000), (1, 96, 2049),
                      (1, 192, 130), (2, 384, 700), (1, 768, 64), (1, 1536, 40), (3, 24, 50001), (1, 80, 97), (2, 16, 4096)]:
        x = (rng.standard_normal((B, T, C)) * 1.5).astype(np.float32)
        al = (0.5 * rng.standard_normal(C)).astype(np.float32)
        be = (0.5 * rng.standard_normal(C)).astype(np.float32)
        xt = torch.from_numpy(x).to(dev)
        alt, bet = torch.from_numpy(al).to(dev), torch.from_numpy(be).to(dev)
        for in_dt, out_dt in ((torch.float32, torch.bfloat16), (torch.bfloat16, torch.bfloat16), (torch.float32, torch.float32)):
            xin = xt.to(in_dt)
            y = ops.activation1d_nlc(xin, alt, bet, out_dt, False)
            torch.cuda.synchronize()
            if B * T * C <= 300000:
                ref = np.transpose(O.activation1d(np.transpose(xin.float().cpu().numpy(), (0, 2, 1)).astype(np.float64), al, be), (0, 2, 1))
            else:
                ref = ops.activation1d_nlc(xin.float(), alt, bet, torch.float32, True).cpu().numpy()
            yn = y.float().cpu().numpy()
            s = snr(ref, yn)
            err = float(np.abs(ref - yn).max())
            # worst rows: where is the error?
            e_t = np.abs(ref - yn).max(axis=(0, 2))
            ok = s > (45.0 if out_dt == torch.bfloat16 else 60.0) and np.isfinite(yn).all()
            bad += not ok
            print(f"B={B} C={C} T={T} {str(in_dt)[6:]}->{str(out_dt)[6:]}: SNR {s:.1f} dB max-abs {err:.2e} "
                  f"worst t={int(e_t.argmax())} {'ok' if ok else 'FAIL'}", flush=True)
    print("CHECK", "FAILED" if bad else "ok", flush=True)

if a.time:
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    rows = []
    for (C, T) in [(768, 3448), (384, 13792), (192, 27584), (96, 55168), (48, 110336), (24, 220672)]:
        B = 16
        for in_dt, out_dt in ((torch.float32, torch.bfloat16), (torch.bfloat16, torch.bfloat16)):
            x = torch.randn(B, T, C, device=dev).to(in_dt)
            al = 0.5 * torch.randn(C, device=dev); be = 0.5 * torch.randn(C, device=dev)
            for _ in range(2):
                y = ops.activation1d_nlc(x, al, be, out_dt, False)
            torch.cuda.synchronize()
            ms = 0.0
            for _ in range(5):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); y = ops.activation1d_nlc(x, al, be, out_dt, False); e1.record()
                torch.cuda.synchronize()
                ms += e0.elapsed_time(e1)
            ms /= 5
            nbytes = B * T * C * (x.element_size() + y.element_size())
            row = dict(C=C, T=T, B=B, io=f"{str(in_dt)[6:]}->{str(out_dt)[6:]}", us=ms * 1e3, GBs=nbytes / ms / 1e6,
                       umma=os.environ.get("BVG_ACT_UMMA", "0"))
            rows.append(row)
            print(json.dumps(row), flush=True)
sys.exit(1 if bad else 0)
