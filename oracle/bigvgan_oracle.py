"""NumPy restatement of the reference BigVGAN-v2 forward (TEST INFRASTRUCTURE, not product).

Every function cites the reference file:line it restates.  Paths are relative to
``/root/reference/indextts/s2mel/modules/bigvgan`` (``BV/``) and
``BV/alias_free_activation`` (``AF/``).

The arithmetic is written from the closed forms (polyphase up-sampler, clamped-index
low-pass, polyphase transposed convolution, per-tap GEMM convolution), i.e. it does
NOT call torch conv kernels; it is pinned against the reference's torch path by
``tests/golden/make_golden.py`` (run in the authoring container, where
``/root/reference`` exists) and ``tests/test_oracle.py`` (runs anywhere, against the
committed ``tests/golden/*.npz``).
"""
from __future__ import annotations

import json
import math
from typing import Dict, Mapping, Optional, Sequence

import numpy as np

__all__ = [
    "DEFAULT_HPARAMS",
    "kaiser_sinc_filter1d",
    "upsample2x",
    "downsample2x",
    "snake_beta",
    "activation1d",
    "conv1d",
    "conv_transpose1d",
    "conv_transpose1d_scatter",
    "fold_weight_norm",
    "get_padding",
    "effective_weights",
    "forward",
    "forward_v1",
    "V1_HPARAMS",
    "forward_time_sharded",
    "make_state_dict",
    "load_hparams",
    "snr_db",
    "RECEPTIVE_HALO_FRAMES",
]

# BV/config.json:1-63 (only the keys the generator reads)
DEFAULT_HPARAMS = {
    "resblock": "1",
    "upsample_rates": [4, 4, 2, 2, 2, 2],
    "upsample_kernel_sizes": [8, 8, 4, 4, 4, 4],
    "upsample_initial_channel": 1536,
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "use_tanh_at_final": False,
    "use_bias_at_final": False,
    "activation": "snakebeta",
    "snake_logscale": True,
    "num_mels": 80,
    "hop_size": 256,
    "sampling_rate": 22050,
}

# Generator section of the v1 model (indextts/BigVGAN/models.py:130-200 reads these keys).  The v1 checkpoint's config file
# is a download and NOT part of the reference tree: the values below are the shape class upstream IndexTTS v1 publishes
# (24 kHz, 1024x up-sampling of the 1024-wide GPT latent, 512-wide speaker embedding); tests use scaled-down versions.
V1_HPARAMS = {
    "resblock": "1",
    "upsample_rates": [4, 4, 4, 4, 2, 2],
    "upsample_kernel_sizes": [8, 8, 4, 4, 4, 4],
    "upsample_initial_channel": 1536,
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "activation": "snakebeta",
    "snake_logscale": True,
    "feat_upsample": False,
    "cond_d_vector_in_each_upsampling_layer": True,
    "speaker_embedding_dim": 512,
    "gpt_dim": 1024,
    "num_mels": 100,
    "sampling_rate": 24000,
}

# one-sided receptive field of the whole generator in mel frames.
RECEPTIVE_HALO_FRAMES = 38


def load_hparams(path: str) -> dict:
    """BV/bigvgan.py:25-28 ``load_hparams_from_json`` (plain dict here)."""
    with open(path) as f:
        return json.load(f)


def get_padding(kernel_size: int, dilation: int = 1) -> int:
    """BV/utils.py:57-58."""
    return int((kernel_size * dilation - dilation) / 2)


# ----------------------------------------------------------------------------------
# Alias-free activation
# ----------------------------------------------------------------------------------
def _i0(x: np.ndarray) -> np.ndarray:
    return np.i0(x)


def kaiser_sinc_filter1d(cutoff: float = 0.25, half_width: float = 0.3, kernel_size: int = 12,
                         dtype=np.float64) -> np.ndarray:
    """AF/torch/filter.py:30-62.  Returns the ``kernel_size`` taps (1-D)."""
    even = kernel_size % 2 == 0
    half_size = kernel_size // 2
    delta_f = 4 * half_width
    A = 2.285 * (half_size - 1) * math.pi * delta_f + 7.95
    if A > 50.0:
        beta = 0.1102 * (A - 8.7)
    elif A >= 21.0:
        beta = 0.5842 * (A - 21) ** 0.4 + 0.07886 * (A - 21.0)
    else:
        beta = 0.0
    # torch.kaiser_window(N, periodic=False, beta) == np.kaiser(N, beta)
    window = np.kaiser(kernel_size, beta)
    if even:
        time = np.arange(-half_size, half_size) + 0.5
    else:
        time = np.arange(kernel_size) - half_size
    if cutoff == 0:
        return np.zeros(kernel_size, dtype=dtype)
    filt = 2 * cutoff * window * np.sinc(2 * cutoff * time)
    filt = filt / filt.sum()
    return filt.astype(dtype)


def _clamped(x: np.ndarray, idx: np.ndarray) -> np.ndarray:
    T = x.shape[-1]
    return x[..., np.clip(idx, 0, T - 1)]


def upsample2x(x: np.ndarray, filt: np.ndarray) -> np.ndarray:
    """AF/torch/resample.py:30-38 (``UpSample1d.forward`` with ratio 2, kernel 12).

    replicate-pad 5/5, zero-stuffed depthwise transposed conv, x2 gain, slice [15:-15].
    Polyphase closed form, ``cl`` = clamp to [0, T-1]:
        u[2i]   = 2 * sum_{m<6} f[2m+1] * x[cl(i+2-m)]
        u[2i+1] = 2 * sum_{m<6} f[2m]   * x[cl(i+3-m)]
    """
    T = x.shape[-1]
    i = np.arange(T)
    even = np.zeros_like(x)
    odd = np.zeros_like(x)
    for m in range(6):
        even = even + filt[2 * m + 1] * _clamped(x, i + 2 - m)
        odd = odd + filt[2 * m] * _clamped(x, i + 3 - m)
    u = np.empty(x.shape[:-1] + (2 * T,), dtype=x.dtype)
    u[..., 0::2] = 2 * even
    u[..., 1::2] = 2 * odd
    return u


def downsample2x(v: np.ndarray, filt: np.ndarray) -> np.ndarray:
    """AF/torch/resample.py:56-58 -> AF/torch/filter.py:94-101 (stride 2, kernel 12).

    replicate-pad 5 left / 6 right of the 2x signal, depthwise stride-2 FIR:
        y[i] = sum_{k<12} g[k] * v[clamp(2i + k - 5, 0, 2T-1)]
    """
    T2 = v.shape[-1]
    T = T2 // 2
    i = np.arange(T)
    y = np.zeros(v.shape[:-1] + (T,), dtype=v.dtype)
    for k in range(12):
        y = y + filt[k] * _clamped(v, 2 * i + k - 5)
    return y


def snake_beta(x: np.ndarray, alpha: np.ndarray, beta: np.ndarray, logscale: bool = True) -> np.ndarray:
    """BV/activations.py:107-120: ``x + 1/(b + 1e-9) * sin(x*a)^2`` per channel ([.., C, T])."""
    a = alpha.astype(x.dtype)[:, None]
    b = beta.astype(x.dtype)[:, None]
    if logscale:
        a = np.exp(a)
        b = np.exp(b)
    return x + (1.0 / (b + x.dtype.type(1e-9))) * np.sin(x * a) ** 2


def activation1d(x: np.ndarray, alpha: np.ndarray, beta: np.ndarray, logscale: bool = True,
                 filt: Optional[np.ndarray] = None) -> np.ndarray:
    """AF/torch/act.py:25-30: down(act(up(x))) on ``[B, C, T]``."""
    if filt is None:
        filt = kaiser_sinc_filter1d(dtype=np.float32).astype(x.dtype)
    else:
        filt = np.asarray(filt).reshape(-1).astype(x.dtype)
    u = upsample2x(x, filt)
    v = snake_beta(u, alpha, beta, logscale)
    return downsample2x(v, filt)


# ----------------------------------------------------------------------------------
# Convolutions
# ----------------------------------------------------------------------------------
def conv1d(x: np.ndarray, w: np.ndarray, bias: Optional[np.ndarray], dilation: int = 1,
           padding: Optional[int] = None) -> np.ndarray:
    """``nn.Conv1d`` as used at BV/bigvgan.py:58-66,75-83,285-287,348-350.

    y[b,co,t] = bias[co] + sum_ci sum_j w[co,ci,j] * x[b,ci,t + j*d - pad], zero outside.
    Evaluated as one GEMM per tap.
    """
    B, Cin, T = x.shape
    Cout, Cin2, k = w.shape
    assert Cin == Cin2
    if padding is None:
        padding = get_padding(k, dilation)
    xp = np.zeros((B, Cin, T + 2 * padding), dtype=x.dtype)
    xp[:, :, padding:padding + T] = x
    Tout = T + 2 * padding - dilation * (k - 1)
    y = np.zeros((B, Cout, Tout), dtype=x.dtype)
    for b in range(B):
        acc = y[b]
        for j in range(k):
            acc += w[:, :, j] @ xp[b, :, j * dilation:j * dilation + Tout]
    if bias is not None:
        y += bias.astype(x.dtype)[None, :, None]
    return y


def conv_transpose1d(x: np.ndarray, w: np.ndarray, bias: Optional[np.ndarray], stride: int,
                     padding: int) -> np.ndarray:
    """``nn.ConvTranspose1d`` at BV/bigvgan.py:305-313, polyphase form.

    y[b,co,u*q+r] = bias[co] + sum_ci sum_{k' = r+p (mod u)} x[b,ci,q + (r+p-k')/u] * w[ci,co,k']
    Output length (Tin-1)*u - 2p + k  (= u*Tin for k = 2u, p = u/2).
    """
    B, Cin, Tin = x.shape
    Cin2, Cout, k = w.shape
    assert Cin == Cin2
    u, p = stride, padding
    Tout = (Tin - 1) * u - 2 * p + k
    y = np.zeros((B, Cout, Tout), dtype=x.dtype)
    for r in range(u):
        n_q = (Tout - r + u - 1) // u
        q = np.arange(n_q)
        for kk in range(k):
            if (r + p - kk) % u != 0:
                continue
            s = (r + p - kk) // u
            n = q + s
            valid = (n >= 0) & (n < Tin)
            if not valid.any():
                continue
            qv = q[valid]
            wk = w[:, :, kk].T  # [Cout, Cin]
            for b in range(B):
                y[b][:, qv * u + r] += wk @ x[b][:, n[valid]]
    if bias is not None:
        y += bias.astype(x.dtype)[None, :, None]
    return y


def conv_transpose1d_scatter(x, w, bias, stride, padding):
    """Definition-form transposed conv (scatter), used only to cross-check the polyphase form."""
    B, Cin, Tin = x.shape
    _, Cout, k = w.shape
    full = np.zeros((B, Cout, (Tin - 1) * stride + k), dtype=x.dtype)
    for n in range(Tin):
        # contribution of input column n to outputs n*stride + kk
        full[:, :, n * stride:n * stride + k] += np.einsum("bi,iok->bok", x[:, :, n], w)
    y = full[:, :, padding:full.shape[-1] - padding]
    if bias is not None:
        y = y + bias.astype(x.dtype)[None, :, None]
    return y


def fold_weight_norm(v: np.ndarray, g: np.ndarray) -> np.ndarray:
    """``torch._weight_norm(v, g, dim=0)`` as removed at BV/bigvgan.py:388-400.

    w = g * v / ||v|| with the norm over every dim except 0 (dim 0 is C_out for Conv1d and
    C_in for ConvTranspose1d).
    """
    norm = np.sqrt((v.astype(np.float64) ** 2).sum(axis=tuple(range(1, v.ndim)), keepdims=True))
    return (v * (g / norm.astype(v.dtype))).astype(v.dtype)


def effective_weights(sd: Mapping[str, np.ndarray], prefix: str, dtype) -> np.ndarray:
    if prefix + ".weight" in sd:
        return np.asarray(sd[prefix + ".weight"]).astype(dtype)
    # the reference folds in the parameters' own dtype (fp32) at remove_weight_norm() time
    return fold_weight_norm(np.asarray(sd[prefix + ".weight_v"]), np.asarray(sd[prefix + ".weight_g"])).astype(dtype)


def _bias(sd, prefix, dtype):
    b = sd.get(prefix + ".bias")
    return None if b is None else np.asarray(b).astype(dtype)


# ----------------------------------------------------------------------------------
# Whole generator
# ----------------------------------------------------------------------------------
def _amp_block1(sd, prefix, x, k, dilations, logscale, dtype, filt):
    """BV/bigvgan.py:132-141 ``AMPBlock1.forward``."""
    for l, d in enumerate(dilations):
        a1, a2 = 2 * l, 2 * l + 1
        xt = activation1d(x, np.asarray(sd[f"{prefix}.activations.{a1}.act.alpha"]),
                          np.asarray(sd[f"{prefix}.activations.{a1}.act.beta"]), logscale, filt)
        xt = conv1d(xt, effective_weights(sd, f"{prefix}.convs1.{l}", dtype),
                    _bias(sd, f"{prefix}.convs1.{l}", dtype), dilation=d)
        xt = activation1d(xt, np.asarray(sd[f"{prefix}.activations.{a2}.act.alpha"]),
                          np.asarray(sd[f"{prefix}.activations.{a2}.act.beta"]), logscale, filt)
        xt = conv1d(xt, effective_weights(sd, f"{prefix}.convs2.{l}", dtype),
                    _bias(sd, f"{prefix}.convs2.{l}", dtype), dilation=1)
        x = xt + x
    return x


def _act(sd, prefix, x, h, dtype, filt):
    """``Activation1d`` of a ``Snake`` (BV/activations.py:54-57: beta := alpha, no ``.act.beta`` key) or ``SnakeBeta``."""
    alpha = np.asarray(sd[f"{prefix}.act.alpha"])
    beta = np.asarray(sd[f"{prefix}.act.beta"]) if h["activation"] == "snakebeta" else alpha
    return activation1d(x, alpha, beta, bool(h["snake_logscale"]), filt)


def _amp_block2(sd, prefix, x, k, dilations, h, dtype, filt):
    """``AMPBlock2.forward`` as upstream wrote it (indextts/BigVGAN/bigvgan.py:231-236, indextts/BigVGAN/models.py:117-123):
    act -> conv(d) -> +x per dilation.  The copy at BV/bigvgan.py:232-236 lacks the ``return x``."""
    for l, d in enumerate(dilations):
        xt = _act(sd, f"{prefix}.activations.{l}", x, h, dtype, filt)
        xt = conv1d(xt, effective_weights(sd, f"{prefix}.convs.{l}", dtype), _bias(sd, f"{prefix}.convs.{l}", dtype),
                    dilation=d)
        x = xt + x
    return x


def _amp_block(sd, prefix, x, k, dilations, h, dtype, filt):
    if h["resblock"] == "2":
        return _amp_block2(sd, prefix, x, k, dilations, h, dtype, filt)
    if h["activation"] == "snakebeta":
        return _amp_block1(sd, prefix, x, k, dilations, bool(h["snake_logscale"]), dtype, filt)
    # AMPBlock1 with Snake activations
    for l, d in enumerate(dilations):
        xt = _act(sd, f"{prefix}.activations.{2 * l}", x, h, dtype, filt)
        xt = conv1d(xt, effective_weights(sd, f"{prefix}.convs1.{l}", dtype), _bias(sd, f"{prefix}.convs1.{l}", dtype),
                    dilation=d)
        xt = _act(sd, f"{prefix}.activations.{2 * l + 1}", xt, h, dtype, filt)
        xt = conv1d(xt, effective_weights(sd, f"{prefix}.convs2.{l}", dtype), _bias(sd, f"{prefix}.convs2.{l}", dtype))
        x = xt + x
    return x


def forward_v1(sd: Mapping[str, np.ndarray], x: np.ndarray, speaker_embedding: np.ndarray, h: Optional[Mapping] = None,
               dtype=np.float32) -> np.ndarray:
    """indextts/BigVGAN/models.py:201-258 ``BigVGAN.forward`` of the v1 generator AFTER the speaker encoder (:202):
    ``x`` ``[B, T, gpt_dim]`` latent, ``speaker_embedding`` ``[B, D]`` -> ``[B, 1, T*prod(rates)]``.
    ``feat_upsample`` (:214-219, a 4x linear interpolation of the latent) must be false, as in the published config."""
    h = dict(V1_HPARAMS if h is None else h)
    assert not h.get("feat_upsample", False)
    filt = kaiser_sinc_filter1d(dtype=np.float32).astype(dtype)
    e = np.asarray(speaker_embedding).astype(dtype)

    def cond(prefix):  # Conv1d(D, C, 1) on the [B, D, 1] embedding (:197-202): [B, C, 1]
        w = np.asarray(sd[prefix + ".weight"]).astype(dtype)[:, :, 0]
        return (e @ w.T + np.asarray(sd[prefix + ".bias"]).astype(dtype))[:, :, None]

    x = np.asarray(x).astype(dtype).transpose(0, 2, 1)  # :221
    x = conv1d(x, effective_weights(sd, "conv_pre", dtype), _bias(sd, "conv_pre", dtype))
    x = x + cond("cond_layer")  # :236
    nk = len(h["resblock_kernel_sizes"])
    for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
        x = conv_transpose1d(x, effective_weights(sd, f"ups.{i}.0", dtype), _bias(sd, f"ups.{i}.0", dtype),
                             stride=u, padding=(k - u) // 2)
        if h.get("cond_d_vector_in_each_upsampling_layer", False):
            x = x + cond(f"conds.{i}")  # :244-245
        xs = None
        for j, (rk, rd) in enumerate(zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"])):
            r = _amp_block(sd, f"resblocks.{i * nk + j}", x, rk, rd, h, dtype, filt)
            xs = r if xs is None else xs + r
        x = xs / dtype(nk)
    x = _act(sd, "activation_post", x, h, dtype, filt)
    x = conv1d(x, effective_weights(sd, "conv_post", dtype), _bias(sd, "conv_post", dtype))
    return np.tanh(x)  # :256


def forward(sd: Mapping[str, np.ndarray], mel: np.ndarray, h: Optional[Mapping] = None,
            dtype=np.float32, taps: Optional[Dict[str, np.ndarray]] = None) -> np.ndarray:
    """BV/bigvgan.py:360-386 ``BigVGAN.forward``: ``[B, num_mels, T] -> [B, 1, T*prod(rates)]``.

    ``taps`` (optional dict) receives named intermediates for per-layer parity tests.
    """
    h = dict(DEFAULT_HPARAMS if h is None else h)
    assert h["resblock"] in ("1", "2") and h["activation"] in ("snake", "snakebeta")
    filt = kaiser_sinc_filter1d(dtype=np.float32).astype(dtype)
    x = np.asarray(mel).astype(dtype)
    x = conv1d(x, effective_weights(sd, "conv_pre", dtype), _bias(sd, "conv_pre", dtype))
    if taps is not None:
        taps["conv_pre"] = x
    nk = len(h["resblock_kernel_sizes"])
    for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
        x = conv_transpose1d(x, effective_weights(sd, f"ups.{i}.0", dtype), _bias(sd, f"ups.{i}.0", dtype),
                             stride=u, padding=(k - u) // 2)
        if taps is not None:
            taps[f"ups.{i}"] = x
        xs = None
        for j, (rk, rd) in enumerate(zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"])):
            r = _amp_block(sd, f"resblocks.{i * nk + j}", x, rk, rd, h, dtype, filt)
            xs = r if xs is None else xs + r
        x = xs / dtype(nk)
        if taps is not None:
            taps[f"stage.{i}"] = x
    x = _act(sd, "activation_post", x, h, dtype, filt)
    x = conv1d(x, effective_weights(sd, "conv_post", dtype), _bias(sd, "conv_post", dtype))
    if taps is not None:
        taps["conv_post"] = x
    if h.get("use_tanh_at_final", True):
        x = np.tanh(x)
    else:
        x = np.clip(x, -1.0, 1.0)
    return x


def forward_time_sharded(sd, mel, n_shards: int, h=None, dtype=np.float32,
                         halo: int = RECEPTIVE_HALO_FRAMES) -> np.ndarray:
    """split the mel along time into ``n_shards`` contiguous frame ranges, give each
    ``halo`` frames of real neighbour context on interior sides, run the ordinary forward, discard the
    ``hop*halo`` samples computed from the halo, concatenate.  True sequence ends keep the reference's own
    padding.  With halo >= 38 this equals the unsharded forward."""
    h = dict(DEFAULT_HPARAMS if h is None else h)
    hop = int(np.prod(h["upsample_rates"]))
    T = mel.shape[-1]
    bounds = [(T * s) // n_shards for s in range(n_shards + 1)]
    outs = []
    for s in range(n_shards):
        lo, hi = bounds[s], bounds[s + 1]
        if hi <= lo:
            continue
        lo_h, hi_h = max(0, lo - halo), min(T, hi + halo)
        w = forward(sd, mel[..., lo_h:hi_h], h, dtype)
        outs.append(w[..., (lo - lo_h) * hop:(lo - lo_h) * hop + (hi - lo) * hop])
    return np.concatenate(outs, axis=-1)


def make_state_dict(h: Optional[Mapping] = None, seed: int = 0, weight_norm: bool = False,
                    alpha_beta_std: float = 0.5, scale: float = 1.0) -> Dict[str, np.ndarray]:
    """Deterministic synthetic state dict with the reference's key names/shapes.

    NOT the reference's initialisation (that needs torch's RNG stream): weights ~ U(-b, b) with the
    Kaiming bound b = 1/sqrt(fan_in) like ``nn.Conv1d``'s default, alpha/beta ~ N(0, alpha_beta_std).
    Used where a test only needs *some* weights shared by oracle and CUDA path.
    """
    h = dict(DEFAULT_HPARAMS if h is None else h)
    rng = np.random.default_rng(seed)
    sd: Dict[str, np.ndarray] = {}
    filt = kaiser_sinc_filter1d(dtype=np.float32).reshape(1, 1, 12)

    def conv(prefix, cout, cin, k, bias=True, transpose=False):
        shape = (cin, cout, k) if transpose else (cout, cin, k)
        fan_in = (cout if transpose else cin) * k
        bound = scale / math.sqrt(fan_in)
        w = rng.uniform(-bound, bound, size=shape).astype(np.float32)
        if weight_norm:
            sd[prefix + ".weight_g"] = np.sqrt((w.astype(np.float64) ** 2).sum(axis=(1, 2), keepdims=True)).astype(
                np.float32) * rng.uniform(0.5, 1.5, size=(shape[0], 1, 1)).astype(np.float32)
            sd[prefix + ".weight_v"] = w
        else:
            sd[prefix + ".weight"] = w
        if bias:
            sd[prefix + ".bias"] = rng.uniform(-bound, bound, size=(cout,)).astype(np.float32)

    def act(prefix, c):
        sd[prefix + ".act.alpha"] = (alpha_beta_std * rng.standard_normal(c)).astype(np.float32)
        if h["activation"] == "snakebeta":
            sd[prefix + ".act.beta"] = (alpha_beta_std * rng.standard_normal(c)).astype(np.float32)
        sd[prefix + ".upsample.filter"] = filt.copy()
        sd[prefix + ".downsample.lowpass.filter"] = filt.copy()

    c0 = h["upsample_initial_channel"]
    v1 = "gpt_dim" in h  # v1 generator (indextts/BigVGAN/models.py): latent input, speaker conditioning, biased conv_post
    conv("conv_pre", c0, h["gpt_dim"] if v1 else h["num_mels"], 7)
    nk = len(h["resblock_kernel_sizes"])
    ch = c0
    for i, (u, k) in enumerate(zip(h["upsample_rates"], h["upsample_kernel_sizes"])):
        conv(f"ups.{i}.0", c0 // 2 ** (i + 1), c0 // 2 ** i, k, transpose=True)
        ch = c0 // 2 ** (i + 1)
        for j, (rk, rd) in enumerate(zip(h["resblock_kernel_sizes"], h["resblock_dilation_sizes"])):
            p = f"resblocks.{i * nk + j}"
            if h["resblock"] == "2":
                for l in range(len(rd)):
                    conv(f"{p}.convs.{l}", ch, ch, rk)
                    act(f"{p}.activations.{l}", ch)
                continue
            for l in range(len(rd)):
                conv(f"{p}.convs1.{l}", ch, ch, rk)
                conv(f"{p}.convs2.{l}", ch, ch, rk)
            for m in range(2 * len(rd)):
                act(f"{p}.activations.{m}", ch)
    act("activation_post", ch)
    conv("conv_post", 1, ch, 7, bias=bool(h.get("use_bias_at_final", True)))
    if v1:  # plain (not weight-normed) 1x1 convs, models.py:197-202
        wn, weight_norm = weight_norm, False
        D = h["speaker_embedding_dim"]
        conv("cond_layer", c0, D, 1)
        if h.get("cond_d_vector_in_each_upsampling_layer", False):
            for i in range(len(h["upsample_rates"])):
                conv(f"conds.{i}", c0 // 2 ** (i + 1), D, 1)
        weight_norm = wn
    return sd


def snr_db(ref: np.ndarray, test: np.ndarray) -> float:
    ref = np.asarray(ref, dtype=np.float64)
    err = np.asarray(test, dtype=np.float64) - ref
    den = float((err ** 2).sum())
    if den == 0.0:
        return float("inf")
    return 10.0 * math.log10(float((ref ** 2).sum()) / den)
