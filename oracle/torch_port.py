"""Functional torch-CPU restatement of the reference forward, op for op (TEST INFRASTRUCTURE, not product).

Where ``bigvgan_oracle.py`` restates the closed-form arithmetic in NumPy, this file restates the reference's own
*operator sequence* -- the same ATen CPU kernels the reference's torch path dispatches to -- so that timing it on the
GPU box's host cores is a faithful stand-in for "the reference's CPU path" (the reference itself is Python under
/root/reference and cannot travel to the box).  Used by ``bench.py`` (``cpu_baseline`` and ``--impl reference``) and
pinned against the golden vectors in ``tests/test_oracle.py``.

Reference lines: BV = indextts/s2mel/modules/bigvgan, AF = BV/alias_free_activation/torch.
"""
from __future__ import annotations

from typing import Mapping

import torch
import torch.nn.functional as F

from . import bigvgan_oracle as O


def _fold(sd: Mapping[str, torch.Tensor], prefix: str) -> torch.Tensor:
    if prefix + ".weight" in sd:
        return sd[prefix + ".weight"]
    return torch._weight_norm(sd[prefix + ".weight_v"], sd[prefix + ".weight_g"], 0)  # BV/bigvgan.py:388-400


def prepare(sd_np: Mapping, dtype=torch.float32):
    """Fold weight norm once (the reference does it at load time, infer_v2.py:263) and convert to tensors."""
    sd = {k: torch.as_tensor(v).to(dtype) for k, v in sd_np.items()}
    out = {}
    for k in list(sd.keys()):
        if k.endswith(".weight_v"):
            p = k[: -len(".weight_v")]
            out[p + ".weight"] = _fold(sd, p)
        elif not k.endswith(".weight_g"):
            out[k] = sd[k]
    return out


def activation1d(x, alpha, beta, filt):
    """AF/act.py:25-30 = UpSample1d (AF/resample.py:30-38) -> SnakeBeta (BV/activations.py:107-120) ->
    DownSample1d/LowPassFilter1d (AF/filter.py:94-101)."""
    C = x.shape[1]
    f = filt.expand(C, -1, -1)
    x = F.pad(x, (5, 5), mode="replicate")
    x = 2 * F.conv_transpose1d(x, f, stride=2, groups=C)
    x = x[..., 15:-15]
    a = torch.exp(alpha)[None, :, None]
    b = torch.exp(beta)[None, :, None]
    x = x + (1.0 / (b + 1e-9)) * torch.pow(torch.sin(x * a), 2)
    x = F.pad(x, (5, 6), mode="replicate")
    return F.conv1d(x, f, stride=2, groups=C)


def forward(sd: Mapping[str, torch.Tensor], mel: torch.Tensor, h=None) -> torch.Tensor:
    """BV/bigvgan.py:360-386 with AMPBlock1 (:132-141); ``sd`` from ``prepare``."""
    h = dict(O.DEFAULT_HPARAMS if h is None else h)
    filt = sd["activation_post.upsample.filter"]
    nk = len(h["resblock_kernel_sizes"])
    with torch.no_grad():
        x = F.conv1d(mel, sd["conv_pre.weight"], sd["conv_pre.bias"], padding=3)
        for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
            x = F.conv_transpose1d(x, sd[f"ups.{i}.0.weight"], sd[f"ups.{i}.0.bias"], stride=u, padding=(k - u) // 2)
            xs = None
            for j, (rk, rd) in enumerate(zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"])):
                p = f"resblocks.{i * nk + j}"
                xr = x
                for l, d in enumerate(rd):
                    xt = activation1d(xr, sd[f"{p}.activations.{2 * l}.act.alpha"], sd[f"{p}.activations.{2 * l}.act.beta"], filt)
                    xt = F.conv1d(xt, sd[f"{p}.convs1.{l}.weight"], sd[f"{p}.convs1.{l}.bias"], dilation=d,
                                  padding=O.get_padding(rk, d))
                    xt = activation1d(xt, sd[f"{p}.activations.{2 * l + 1}.act.alpha"],
                                      sd[f"{p}.activations.{2 * l + 1}.act.beta"], filt)
                    xt = F.conv1d(xt, sd[f"{p}.convs2.{l}.weight"], sd[f"{p}.convs2.{l}.bias"], padding=O.get_padding(rk, 1))
                    xr = xt + xr
                xs = xr if xs is None else xs + xr
            x = xs / nk
        x = activation1d(x, sd["activation_post.act.alpha"], sd["activation_post.act.beta"], filt)
        x = F.conv1d(x, sd["conv_post.weight"], sd.get("conv_post.bias"), padding=3)
        if h.get("use_tanh_at_final", True):
            return torch.tanh(x)
        return torch.clamp(x, min=-1.0, max=1.0)
