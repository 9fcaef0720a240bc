"""Shared helpers for the parity tests (test infrastructure)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import bigvgan_oracle as O  # noqa: E402

MINI_H = dict(O.DEFAULT_HPARAMS, upsample_initial_channel=512)


def build_model(h: dict, sd_np: dict, device="cuda", precision="fp32"):
    """Our BigVGAN with the given NumPy state dict (plain or weight-norm form) on `device`."""
    from index_tts_b200 import bigvgan as B

    m = B.BigVGAN(B.AttrDict(dict(h)), use_cuda_kernel=True)
    if not any(k.endswith("weight_v") for k in sd_np):
        m.remove_weight_norm()
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd_np.items()}, strict=True)
    m = m.to(device).eval()
    m.set_precision(precision)
    return m


def snr_db(ref, test):
    return O.snr_db(np.asarray(ref), np.asarray(test))


def rel_err(ref, test):
    ref = np.asarray(ref, dtype=np.float64)
    return float(np.abs(np.asarray(test, dtype=np.float64) - ref).max() / (np.abs(ref).max() + 1e-30))
