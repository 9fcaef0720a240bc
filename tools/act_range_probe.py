import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
from oracle import bigvgan_oracle as O
from index_tts_b200 import ops
rng = np.random.default_rng(5)
for amax, xs in ((1.0, 1.5), (2.0, 3.0), (3.0, 4.0), (4.0, 8.0)):
    C, T = 48, 4096
    x = (rng.standard_normal((2, C, T)) * xs).astype(np.float32)
    al = rng.uniform(-amax, amax, C).astype(np.float32)
    be = rng.uniform(-amax, amax, C).astype(np.float32)
    ref = O.activation1d(x.astype(np.float64), al, be)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    yf = ops.activation1d(t(x), t(al), t(be), False).cpu().numpy()
    yp = ops.activation1d(t(x), t(al), t(be), True).cpu().numpy()
    print(f"|alpha_log|,|beta_log|<={amax} x*{xs}: max|arg|={float(np.abs(x).max()*np.exp(amax)*2):.0f} rad  fast: max-abs {np.abs(yf-ref).max():.2e} SNR {O.snr_db(ref,yf):.1f} dB   precise: max-abs {np.abs(yp-ref).max():.2e} SNR {O.snr_db(ref,yp):.1f} dB  (|ref|max {np.abs(ref).max():.1f})")
