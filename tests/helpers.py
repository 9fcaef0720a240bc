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
# scaled-down v1 generator (indextts/BigVGAN/models.py): 1024x up-sampling, conditioning in every stage
V1_MINI_H = dict(O.V1_HPARAMS, upsample_initial_channel=512, gpt_dim=64, speaker_embedding_dim=32)


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


def build_seeded_ecapa(ecapa_module, seed=77):
    """OUR ECAPA-TDNN with torch-seeded weights and non-trivial BatchNorm statistics (shared with tests/test_v1.py)."""
    torch.manual_seed(seed)
    m = ecapa_module.ECAPA_TDNN(100, lin_neurons=48)
    g = torch.Generator().manual_seed(seed + 1)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm1d):
            mod.running_mean.copy_(0.1 * torch.randn(mod.num_features, generator=g))
            mod.running_var.copy_(1.0 + 0.2 * torch.rand(mod.num_features, generator=g))
            mod.weight.data.copy_(1.0 + 0.1 * torch.randn(mod.num_features, generator=g))
            mod.bias.data.copy_(0.1 * torch.randn(mod.num_features, generator=g))
    return m.eval()
