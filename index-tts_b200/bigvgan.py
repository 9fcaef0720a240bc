"""BigVGAN-v2 generator with the reference's Python surface and a B200-native forward.

Mirrors ``indextts/s2mel/modules/bigvgan/bigvgan.py`` of the reference: ``load_hparams_from_json`` (:25-28),
``AMPBlock1`` (:31-147), ``AMPBlock2`` (:150-240), ``BigVGAN`` (:243-492) -- same constructor arguments, same
module tree, same 783 / 667 state-dict keys, same RNG consumption during construction (so a seeded construction
yields the reference's weights bit for bit), ``remove_weight_norm()``, ``from_pretrained`` / ``_save_pretrained``.

What differs is everything below ``forward``: the module tree only *stores* parameters.  ``BigVGAN.forward`` folds
weight-norm and packs the weights once (tap-major bf16 for the tcgen05 implicit-GEMM kernels, polyphase for the
transposed convs) and then makes ONE call into ``libbvg.so`` (``bvg_forward``), which launches the whole generator
on the caller's current CUDA stream.  ``use_cuda_kernel`` is accepted and stored but selects nothing: the sm_100a
kernels are the only implementation, and a CPU tensor or a non-sm_100 device raises.
"""
from __future__ import annotations

import collections
import ctypes
import json
import os
import threading
import weakref
from pathlib import Path
from typing import Dict, List, Optional, Union

import torch
import torch.nn as nn
from torch.nn.utils import remove_weight_norm, weight_norm

try:  # the reference requires huggingface_hub (bigvgan.py:22); keep working without it for local use
    from huggingface_hub import PyTorchModelHubMixin, hf_hub_download
except Exception:  # pragma: no cover
    class PyTorchModelHubMixin:  # type: ignore
        def __init_subclass__(cls, **kwargs):
            super().__init_subclass__()

    hf_hub_download = None

from . import _lib, activations, ops
from .alias_free_activation.torch.act import Activation1d as TorchActivation1d
from .env import AttrDict
from .utils import get_padding, init_weights

DEFAULT_PRECISION = os.environ.get("BVG_PRECISION", "bf16")


def load_hparams_from_json(path) -> AttrDict:
    with open(path) as f:
        data = f.read()
    return AttrDict(json.loads(data))


# ---------------------------------------------------------------------------------------------------------------
# parameter-holding conv modules whose forward is the libbvg kernel
# ---------------------------------------------------------------------------------------------------------------
def _packed_cache(module, key, build):
    """Cache a derived tensor on ``module`` keyed by the versions of the tensors it was derived from."""
    cache = module.__dict__.setdefault("_bvg_cache", {})
    hit = cache.get(key[0])
    if hit is not None and hit[0] == key:
        return hit[1]
    val = build()
    cache[key[0]] = (key, val)
    return val


def _weight_sources(m):
    if hasattr(m, "weight_v"):
        return (m.weight_v, m.weight_g)
    return (m.weight,)


def _effective_weight(m) -> torch.Tensor:
    """Folded fp32 weight of a (possibly weight-normed) conv module, computed by ``bvg_fold_weight_norm``."""
    if hasattr(m, "weight_v"):
        return ops.fold_weight_norm(m.weight_v, m.weight_g)
    return m.weight.detach().float().contiguous()


def _fingerprint(tensors):
    return tuple((t.data_ptr(), t._version, t.device.index) for t in tensors)


def _to_operand(x_ncl: torch.Tensor, prec: int) -> torch.Tensor:
    """``[B, C, T]`` tensor -> the channels-last conv operand of precision ``prec`` (bf16, fp32 or the bf16x3 split)."""
    if prec == _lib.BVG_F32_SPLIT_TC:
        return ops.split_bf16x3(ops.ncl_to_nlc(x_ncl, torch.float32))
    return ops.ncl_to_nlc(x_ncl, torch.bfloat16 if prec == _lib.BVG_BF16_TC else torch.float32)


class Conv1d(nn.Conv1d):
    """``nn.Conv1d`` as a parameter container; ``forward`` runs the libbvg implicit-GEMM kernel ([B,C,T] in/out).
    Only what the generator uses is supported: stride 1, zeros 'same' padding, groups 1."""

    bvg_precision = None

    def packed(self, precision: int) -> torch.Tensor:
        src = _weight_sources(self)
        return _packed_cache(self, ("w", precision) + _fingerprint(src),
                             lambda: ops.pack_conv1d_weight(_effective_weight(self), precision))

    def forward(self, x):
        k, d = self.kernel_size[0], self.dilation[0]
        if self.stride[0] != 1 or self.groups != 1 or self.padding[0] != get_padding(k, d) or k % 2 == 0:
            raise NotImplementedError("libbvg Conv1d: stride 1, groups 1, odd kernel and 'same' zero padding only")
        prec = ops.precision_code(self.bvg_precision or DEFAULT_PRECISION)
        if self.out_channels % 8 != 0:  # conv_post (C -> 1): the dedicated tail kernel without the clamp is not exposed
            raise NotImplementedError("libbvg Conv1d: out_channels must be a multiple of 8 (conv_post is fused into "
                                      "BigVGAN.forward)")
        xn = _to_operand(x.float() if x.dtype not in (torch.float32, torch.bfloat16) else x, prec)
        bias = None if self.bias is None else self.bias.detach().float().contiguous()
        y = ops.conv1d_nlc(xn, self.packed(prec), bias, d, prec)
        return ops.nlc_to_ncl(y, x.dtype if x.dtype in (torch.float32, torch.bfloat16) else torch.float32)


class ConvTranspose1d(nn.ConvTranspose1d):
    """``nn.ConvTranspose1d`` as a parameter container; ``forward`` runs the polyphase libbvg kernel."""

    bvg_precision = None

    def packed(self, precision: int) -> torch.Tensor:
        src = _weight_sources(self)
        return _packed_cache(self, ("w", precision) + _fingerprint(src),
                             lambda: ops.pack_convtranspose1d_weight(_effective_weight(self), self.stride[0], precision))

    def forward(self, x, output_size=None):
        K, u = self.kernel_size[0], self.stride[0]
        if self.padding[0] != (K - u) // 2 or K % u != 0 or (K - u) % 2 != 0 or self.groups != 1 or \
                self.output_padding[0] != 0 or self.dilation[0] != 1:
            raise NotImplementedError("libbvg ConvTranspose1d: K % stride == 0 and padding (K-stride)/2 only")
        prec = ops.precision_code(self.bvg_precision or DEFAULT_PRECISION)
        xn = _to_operand(x, prec)
        bias = None if self.bias is None else self.bias.detach().float().contiguous()
        y = ops.convtranspose1d_nlc(xn, self.packed(prec), bias, K, u, prec)
        return ops.nlc_to_ncl(y, x.dtype if x.dtype in (torch.float32, torch.bfloat16) else torch.float32)


def _select_activation1d(h):
    # reference bigvgan.py:94-102: lazy import of the cuda flavour when requested.  Both flavours are the same
    # fused kernel here; the import keeps `use_cuda_kernel=True` construction paths identical.
    if h.get("use_cuda_kernel", False):
        from .alias_free_activation.cuda.activation1d import Activation1d as CudaActivation1d
        return CudaActivation1d
    return TorchActivation1d


def _make_activations(h, channels, n, activation, Activation1d):
    if activation == "snake":
        return nn.ModuleList([Activation1d(activation=activations.Snake(channels, alpha_logscale=h.snake_logscale))
                              for _ in range(n)])
    if activation == "snakebeta":
        return nn.ModuleList([Activation1d(activation=activations.SnakeBeta(channels, alpha_logscale=h.snake_logscale))
                              for _ in range(n)])
    raise NotImplementedError(
        "activation incorrectly specified. check the config file and look for 'activation'.")


def _block_precision(block) -> int:
    return ops.precision_code(getattr(block, "bvg_precision", None) or DEFAULT_PRECISION)


class AMPBlock1(torch.nn.Module):
    """Reference bigvgan.py:31-147.  ``forward`` (:132-141) runs act -> conv(d) -> act -> conv(1) -> +x three times
    through the libbvg kernels in channels-last layout; the residual add is fused into the second conv's epilogue."""

    def __init__(self, h: AttrDict, channels: int, kernel_size: int = 3, dilation: tuple = (1, 3, 5),
                 activation: str = None):
        super().__init__()
        self.h = h
        self.convs1 = nn.ModuleList([
            weight_norm(Conv1d(channels, channels, kernel_size, stride=1, dilation=d,
                               padding=get_padding(kernel_size, d))) for d in dilation])
        self.convs1.apply(init_weights)
        self.convs2 = nn.ModuleList([
            weight_norm(Conv1d(channels, channels, kernel_size, stride=1, dilation=1,
                               padding=get_padding(kernel_size, 1))) for _ in range(len(dilation))])
        self.convs2.apply(init_weights)
        self.num_layers = len(self.convs1) + len(self.convs2)
        self.activations = _make_activations(h, channels, self.num_layers, activation, _select_activation1d(self.h))

    def forward(self, x):
        prec = _block_precision(self)
        tc = prec == _lib.BVG_BF16_TC
        op_dtype = ops.operand_dtype(prec)
        xr = ops.ncl_to_nlc(x, torch.float32)  # fp32 residual stream
        acts1, acts2 = self.activations[::2], self.activations[1::2]
        for c1, c2, a1, a2 in zip(self.convs1, self.convs2, acts1, acts2):
            al1, be1 = (t.float().contiguous() for t in a1.log_params())
            al2, be2 = (t.float().contiguous() for t in a2.log_params())
            b1 = None if c1.bias is None else c1.bias.detach().float().contiguous()
            b2 = None if c2.bias is None else c2.bias.detach().float().contiguous()
            xt = ops.activation1d_nlc(xr, al1, be1, op_dtype, not tc)
            xt = ops.conv1d_nlc(xt, c1.packed(prec), b1, c1.dilation[0], prec, out_f32=not tc, out_bf16=tc)
            xt = ops.activation1d_nlc(xt, al2, be2, op_dtype, not tc)
            xr = ops.conv1d_nlc(xt, c2.packed(prec), b2, 1, prec, residual=xr)
        return ops.nlc_to_ncl(xr, x.dtype if x.dtype in (torch.float32, torch.bfloat16) else torch.float32)

    def remove_weight_norm(self):
        for l in self.convs1:
            remove_weight_norm(l)
        for l in self.convs2:
            remove_weight_norm(l)


class AMPBlock2(torch.nn.Module):
    """Reference bigvgan.py:150-240 (one conv per dilation).  Not used by ``bigvgan_v2_22khz_80band_256x``
    (config.json:2 ``resblock "1"``).  NOTE: the reference copy's ``forward`` lacks ``return x`` (:232-236) and so
    returns ``None``; this implementation returns the tensor like upstream BigVGAN."""

    def __init__(self, h: AttrDict, channels: int, kernel_size: int = 3, dilation: tuple = (1, 3, 5),
                 activation: str = None):
        super().__init__()
        self.h = h
        self.convs = nn.ModuleList([
            weight_norm(Conv1d(channels, channels, kernel_size, stride=1, dilation=d,
                               padding=get_padding(kernel_size, d))) for d in dilation])
        self.convs.apply(init_weights)
        self.num_layers = len(self.convs)
        self.activations = _make_activations(h, channels, self.num_layers, activation, _select_activation1d(self.h))

    def forward(self, x):
        prec = _block_precision(self)
        tc = prec == _lib.BVG_BF16_TC
        op_dtype = ops.operand_dtype(prec)
        xr = ops.ncl_to_nlc(x, torch.float32)
        for c, a in zip(self.convs, self.activations):
            al, be = (t.float().contiguous() for t in a.log_params())
            b = None if c.bias is None else c.bias.detach().float().contiguous()
            xt = ops.activation1d_nlc(xr, al, be, op_dtype, not tc)
            xr = ops.conv1d_nlc(xt, c.packed(prec), b, c.dilation[0], prec, residual=xr)
        return ops.nlc_to_ncl(xr, x.dtype if x.dtype in (torch.float32, torch.bfloat16) else torch.float32)

    def remove_weight_norm(self):
        for l in self.convs:
            remove_weight_norm(l)


# ---------------------------------------------------------------------------------------------------------------
# whole-generator launch plan
# ---------------------------------------------------------------------------------------------------------------
_PLANS: "weakref.WeakValueDictionary[int, _PackedGenerator]" = weakref.WeakValueDictionary()
_CAPTURE_LOCK = threading.Lock()
_next_plan_id = [1]


def fold_conv1d_weight(w: torch.Tensor, dilation: int, fold: int) -> torch.Tensor:
    """Time-folded form of a 'same'-padded Conv1d weight (include/bvg.h, ``bvg_conv_desc::fold``).

    ``w`` is ``[Cout, Cin, k]`` (k odd).  Viewing ``[B, T, C]`` as ``[B, T/fold, fold*C]``, output phase ``po`` of folded
    row ``t'`` is original row ``fold*t' + po``; tap ``j`` reads original row ``fold*t' + po + (j - (k-1)/2)*dilation``,
    i.e. folded row ``t' + floor(q / fold)`` phase ``q mod fold`` with ``q = po + (j - (k-1)/2)*dilation``.  Returns
    ``[fold*Cout, fold*Cin, taps']`` with ``taps' = 2*ceil(((k-1)/2*dilation) / fold) + 1`` for a dilation-1 conv."""
    cout, cin, k = w.shape
    reach = (k - 1) // 2 * dilation
    pad = -(-reach // fold)
    taps = 2 * pad + 1
    out = torch.zeros(fold * cout, fold * cin, taps, dtype=w.dtype, device=w.device)
    for po in range(fold):
        for j in range(k):
            q = po + (j - (k - 1) // 2) * dilation
            tt, pi = q // fold, q % fold  # Python floor division / non-negative modulo
            out[po * cout:(po + 1) * cout, pi * cin:(pi + 1) * cin, tt + pad] = w[:, :, j]
    return out


# Time folding of the AMP-block convs of small-channel stages (include/bvg.h, bvg_conv_desc::fold): stages with at most
# FOLD_MAX_CHANNELS channels are folded by FOLD_FACTOR (C = 24 -> 96), wider ones up to C = 64 by 2 (C = 48 -> 96), when
# the folded conv stays within _fold_max_taps() taps.  The folded GEMM does fold*taps'/k times the FLOPs on an otherwise
# idle tensor pipe and gets 4x / 2x fewer, wider tiles in exchange.  BVG_FOLD=0 switches it off.
FOLD_MAX_CHANNELS = 32
FOLD_FACTOR = 4
FOLD_MAX_TAPS = 5


def _fold_max_taps(r: int, k: int) -> int:
    """Measured per layer (gpurun r01, tools/layer_ab.py): folding by 4 pays up to 9 folded taps (the folded layer is a
    C = 96 conv, weight-streaming in pair mode beyond 4 taps); folding by 2 pays up to 7 taps for k >= 7 but loses for
    the k = 3, d = 5 layer.  More than 9 taps exceed the shared-memory plan of the conv kernel."""
    if r >= 4:
        return 9
    return 7 if k >= 7 else FOLD_MAX_TAPS


class _PackedGenerator:
    """Folded + packed weights of one BigVGAN on one device in one precision, and the C descriptor pointing at them."""

    def __init__(self, model: "BigVGAN", device: torch.device, precision: int):
        h = model.h
        self.device = device
        self.precision = precision
        self.keep = []  # tensors the descriptor points into
        self.hop = 1
        for u in h.upsample_rates:
            self.hop *= int(u)
        d = _lib.ModelDesc()
        d.num_mels = int(model.conv_pre.in_channels)  # h.num_mels here; h.gpt_dim for the v1 generator (v1.py)
        d.upsample_initial_channel = int(h.upsample_initial_channel)
        d.num_stages = model.num_upsamples
        d.num_kernels = model.num_kernels
        # dilations per AMP block, read off the modules (the v1 AMPBlock2 keeps two whatever the config lists, v1.py)
        dil = {len(rb.convs if isinstance(rb, AMPBlock2) else rb.convs1) for rb in model.resblocks}
        if len(dil) != 1:
            raise NotImplementedError("resblock_dilation_sizes must have the same length for every kernel size")
        d.num_dilations = dil.pop()
        if d.num_stages > _lib.MAX_STAGES or d.num_kernels > _lib.MAX_KERNELS or d.num_dilations > _lib.MAX_DILATIONS:
            raise NotImplementedError("generator larger than the libbvg descriptor limits")
        for i, (u, k) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
            d.upsample_rates[i] = int(u)
            d.upsample_kernel_sizes[i] = int(k)
        d.use_tanh_at_final = int(bool(model.use_tanh_at_final))
        d.precision = precision

        def f32(t):
            t = t.detach().to(device=device, dtype=torch.float32).contiguous()
            self.keep.append(t)
            return t

        fold_ok = precision == _lib.BVG_BF16_TC and os.environ.get("BVG_FOLD", "1") != "0"

        def conv(dst, m, transposed=False, time_factor=0):
            dst.cin, dst.cout = int(m.in_channels), int(m.out_channels)
            dst.k = int(m.kernel_size[0])
            dst.dilation = int(m.dilation[0])
            dst.fold = 0
            r = FOLD_FACTOR if dst.cin <= FOLD_MAX_CHANNELS else 2  # C = 24 -> 96 by 4, C = 48 -> 96 by 2
            if (fold_ok and not transposed and time_factor % r == 0 and dst.cin == dst.cout
                    and dst.cin * r <= 128 and dst.k % 2 == 1
                    and 2 * -(-((dst.k - 1) // 2 * dst.dilation) // r) + 1 <= _fold_max_taps(r, dst.k)):
                wf = fold_conv1d_weight(_effective_weight(m).detach().to(device=device, dtype=torch.float32),
                                        dst.dilation, r)
                w = ops.pack_conv1d_weight(wf, precision)
                self.keep.append(w)
                dst.w = w.data_ptr()
                dst.bias = f32(m.bias.detach().repeat(r)).data_ptr() if m.bias is not None else None
                dst.fold = r
                return
            w = m.packed(precision)
            self.keep.append(w)
            dst.w = w.data_ptr()
            dst.bias = f32(m.bias).data_ptr() if m.bias is not None else None

        def act(dst, a):
            al, be = a.log_params()
            dst.alpha_log = f32(al).data_ptr()
            dst.beta_log = f32(be).data_ptr()

        conv(d.conv_pre, model.conv_pre)
        time_factor = 1  # T of the stage = time_factor * T_mel: folding by r needs r | time_factor
        for i in range(model.num_upsamples):
            if len(model.ups[i]) != 1:
                raise NotImplementedError("one transposed conv per upsample stage expected")
            conv(d.ups[i], model.ups[i][0], transposed=True)
            time_factor *= int(h.upsample_rates[i])
            for j in range(model.num_kernels):
                rb = model.resblocks[i * model.num_kernels + j]
                kind = 2 if isinstance(rb, AMPBlock2) else (1 if isinstance(rb, AMPBlock1) else 0)
                if kind == 0 or d.resblock not in (0, kind):
                    raise NotImplementedError("the fused generator plan needs all resblocks to be AMPBlock1 or all AMPBlock2")
                d.resblock = kind
                if kind == 1:
                    for l in range(d.num_dilations):
                        conv(d.convs1[i][j][l], rb.convs1[l], time_factor=time_factor)
                        conv(d.convs2[i][j][l], rb.convs2[l], time_factor=time_factor)
                    for a in range(2 * d.num_dilations):
                        act(d.acts[i][j][a], rb.activations[a])
                else:  # AMPBlock2: one conv and one activation per dilation (bvg.h, bvg_model_desc::resblock)
                    for l in range(d.num_dilations):
                        conv(d.convs1[i][j][l], rb.convs[l], time_factor=time_factor)
                        act(d.acts[i][j][l], rb.activations[l])
        act(d.act_post, model.activation_post)
        cp = model.conv_post
        w_post = f32(_effective_weight(cp))
        d.conv_post.w = w_post.data_ptr()
        d.conv_post.bias = f32(cp.bias).data_ptr() if cp.bias is not None else None
        d.conv_post.cin, d.conv_post.cout = int(cp.in_channels), int(cp.out_channels)
        d.conv_post.k, d.conv_post.dilation = int(cp.kernel_size[0]), 1
        self.desc = d
        self.graph_max_frames = int(getattr(model, "_bvg_graph_max_frames", 0) or 0)
        self.plan_id = _next_plan_id[0]
        _next_plan_id[0] += 1
        _PLANS[self.plan_id] = self

    def run_cond(self, x_nlc: torch.Tensor, conds) -> torch.Tensor:
        """v1 generator (v1.py): ``x_nlc`` fp32 ``[B, T, gpt_dim]``, ``conds`` = ``num_stages + 1`` fp32 ``[B, C_i]``
        tensors (or ``None`` entries) added after ``conv_pre`` / each transposed conv (``bvg_forward_cond``)."""
        lib = _lib.load()
        B, T, _ = x_nlc.shape
        wav = torch.empty((B, 1, T * self.hop), dtype=torch.float32, device=x_nlc.device)
        if B == 0 or T == 0:
            return wav
        x_nlc = x_nlc.contiguous()
        n = self.desc.num_stages + 1
        if len(conds) != n:
            raise _lib.BvgError(f"run_cond: expected {n} conditioning vectors, got {len(conds)}")
        conds = [None if c is None else c.detach().to(device=x_nlc.device, dtype=torch.float32).contiguous() for c in conds]
        arr = (ctypes.c_void_p * n)(*[None if c is None else c.data_ptr() for c in conds])
        ws_bytes = lib.bvg_workspace_bytes(ctypes.byref(self.desc), B, T)
        if ws_bytes == 0:
            _lib.check(-1, "bvg_workspace_bytes")
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x_nlc.device)
        with torch.cuda.device(x_nlc.device):
            _lib.check(lib.bvg_forward_cond(ctypes.byref(self.desc), x_nlc.data_ptr(), 1, 0, 0, arr, wav.data_ptr(), B, T,
                                            ws.data_ptr(), ws_bytes, torch.cuda.current_stream().cuda_stream),
                       "bvg_forward_cond")
        return wav

    # ---- CUDA-graph replay for small, launch-bound calls (one 5 s utterance is 214 launches in ~2 ms) -----------------
    # Off unless BigVGAN.set_cuda_graphs(max_frames) / $BVG_GRAPH_MAX_FRAMES says otherwise.  A call with B*T <= max_frames
    # is captured once per (B, T, output dtype) into a graph with static input / output / workspace buffers and replayed
    # afterwards: copy in, one graph launch, copy out.  The cache is per host thread (two server threads never share static
    # buffers); an event orders reuse of the buffers across streams of one thread.  Results are bit-identical to the eager
    # plan (same kernels, same order).
    GRAPH_CACHE_ENTRIES = 4

    def _graphs(self):
        tl = self.__dict__.setdefault("_tl", threading.local())
        if not hasattr(tl, "graphs"):
            tl.graphs = collections.OrderedDict()
        return tl.graphs

    def _run_graphed(self, mel: torch.Tensor, pcm16: bool) -> torch.Tensor:
        cache = self._graphs()
        key = (tuple(mel.shape), bool(pcm16), mel.device)
        ent = cache.get(key)
        cur = torch.cuda.current_stream(mel.device)
        if ent is None:
            static_in = torch.empty(tuple(mel.shape), dtype=torch.float32, device=mel.device)
            static_in.copy_(mel)
            self._run_eager(static_in, pcm16)  # warm-up outside capture (lazy CUDA state, aux streams of the library)
            g = torch.cuda.CUDAGraph()
            # One capture at a time in the process (_CAPTURE_LOCK): two simultaneous captures were seen to invalidate each
            # other.  thread_local error mode: other host threads keep launching / allocating while this one captures
            # (serve_tars.py's two workers); the default global mode would fail THEIR calls with cudaErrorIllegalState.
            # Own capture stream: torch.cuda.graph's default one is shared by all threads of the process.
            with _CAPTURE_LOCK:
                cur.synchronize()
                with torch.cuda.graph(g, stream=torch.cuda.Stream(mel.device), capture_error_mode="thread_local"):
                    static_out = self._run_eager(static_in, pcm16)
            ent = [g, static_in, static_out, None]
            cache[key] = ent
            while len(cache) > self.GRAPH_CACHE_ENTRIES:
                cache.popitem(last=False)
        else:
            cache.move_to_end(key)
            if ent[3] is not None:
                cur.wait_event(ent[3])
        g, static_in, static_out, _ = ent
        static_in.copy_(mel)
        g.replay()
        out = static_out.clone()
        ev = torch.cuda.Event()
        ev.record(cur)
        ent[3] = ev
        return out

    def run(self, mel: torch.Tensor, pcm16: bool = False) -> torch.Tensor:
        B, _, T = mel.shape
        if (self.graph_max_frames > 0 and 0 < B * T <= self.graph_max_frames
                and not torch.cuda.is_current_stream_capturing()):
            with torch.cuda.device(mel.device):
                return self._run_graphed(mel, pcm16)
        return self._run_eager(mel, pcm16)

    def _run_eager(self, mel: torch.Tensor, pcm16: bool = False) -> torch.Tensor:
        lib = _lib.load()
        B, _, T = mel.shape
        wav = torch.empty((B, 1, T * self.hop), dtype=torch.int16 if pcm16 else torch.float32, device=mel.device)
        if B == 0 or T == 0:
            return wav
        if mel.stride(2) != 1:
            mel = mel.contiguous()
        ws_bytes = lib.bvg_workspace_bytes(ctypes.byref(self.desc), B, T)
        if ws_bytes == 0:
            _lib.check(-1, "bvg_workspace_bytes")
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=mel.device)  # per call: safe across host threads
        with torch.cuda.device(mel.device):
            fn = lib.bvg_forward_pcm16 if pcm16 else lib.bvg_forward
            _lib.check(fn(ctypes.byref(self.desc), mel.data_ptr(), mel.stride(0), mel.stride(1), wav.data_ptr(), B, T,
                          ws.data_ptr(), ws_bytes, torch.cuda.current_stream().cuda_stream),
                       "bvg_forward_pcm16" if pcm16 else "bvg_forward")
        return wav


def profile_forward(model: "BigVGAN", mel: torch.Tensor):
    """Measurement helper (bench.py): one forward with a CUDA event pair around every kernel launch
    (``bvg_forward_profile``).  Returns ``(wav, records)``; each record is a dict with the launch kind, layer shape,
    algorithmic flops/bytes and measured milliseconds."""
    lib = _lib.load()
    plan = model._plan(mel.device)
    B, _, T = mel.shape
    wav = torch.empty((B, 1, T * plan.hop), dtype=torch.float32, device=mel.device)
    ws_bytes = lib.bvg_workspace_bytes(ctypes.byref(plan.desc), B, T)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=mel.device)
    recs = (_lib.LaunchRecord * 1024)()
    n = ctypes.c_int(0)
    mel = mel.float().contiguous()
    with torch.cuda.device(mel.device):
        _lib.check(lib.bvg_forward_profile(ctypes.byref(plan.desc), mel.data_ptr(), mel.stride(0), mel.stride(1),
                                           wav.data_ptr(), B, T, ws.data_ptr(), ws_bytes,
                                           torch.cuda.current_stream().cuda_stream, recs, 1024, ctypes.byref(n)),
                   "bvg_forward_profile")
    out = []
    for r in recs[:min(n.value, 1024)]:
        out.append(dict(kind=_lib.KIND_NAMES.get(r.kind, str(r.kind)), stage=r.stage, cin=r.cin, cout=r.cout, k=r.k,
                        dilation=r.dilation, rows=r.rows, flops=r.flops, bytes=r.bytes, ms=r.ms))
    return wav, out


def profile_forward_cond(plan: _PackedGenerator, x_nlc: torch.Tensor, conds):
    """``profile_forward`` for the v1 generator (``bvg_forward_cond_profile``): ``x_nlc`` fp32 ``[B, T, gpt_dim]``,
    ``conds`` as in ``_PackedGenerator.run_cond``.  Returns ``(wav, records)``."""
    lib = _lib.load()
    B, T, _ = x_nlc.shape
    x_nlc = x_nlc.float().contiguous()
    wav = torch.empty((B, 1, T * plan.hop), dtype=torch.float32, device=x_nlc.device)
    n_c = plan.desc.num_stages + 1
    conds = [None if c is None or c.numel() == 0 else c.detach().float().contiguous() for c in conds]
    arr = (ctypes.c_void_p * n_c)(*[None if c is None else c.data_ptr() for c in conds])
    ws_bytes = lib.bvg_workspace_bytes(ctypes.byref(plan.desc), B, T)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x_nlc.device)
    recs = (_lib.LaunchRecord * 1024)()
    n = ctypes.c_int(0)
    with torch.cuda.device(x_nlc.device):
        _lib.check(lib.bvg_forward_cond_profile(ctypes.byref(plan.desc), x_nlc.data_ptr(), 1, 0, 0, arr, wav.data_ptr(), B, T,
                                                ws.data_ptr(), ws_bytes, torch.cuda.current_stream().cuda_stream, recs,
                                                1024, ctypes.byref(n)), "bvg_forward_cond_profile")
    out = [dict(kind=_lib.KIND_NAMES.get(r.kind, str(r.kind)), stage=r.stage, cin=r.cin, cout=r.cout, k=r.k,
                dilation=r.dilation, rows=r.rows, flops=r.flops, bytes=r.bytes, ms=r.ms) for r in recs[:min(n.value, 1024)]]
    return wav, out


@torch.library.custom_op("bvg::vocoder_forward", mutates_args=())
def _vocoder_forward(mel: torch.Tensor, plan_id: int, hop: int) -> torch.Tensor:
    plan = _PLANS.get(plan_id)
    if plan is None:
        raise _lib.BvgError(f"bvg::vocoder_forward: plan {plan_id} no longer exists")
    return plan.run(mel)


@_vocoder_forward.register_fake
def _(mel, plan_id, hop):
    return mel.new_empty((mel.shape[0], 1, mel.shape[2] * hop), dtype=torch.float32)


@torch.library.custom_op("bvg::vocoder_forward_cond", mutates_args=())
def _vocoder_forward_cond(x_nlc: torch.Tensor, conds: List[torch.Tensor], plan_id: int, hop: int) -> torch.Tensor:
    plan = _PLANS.get(plan_id)
    if plan is None:
        raise _lib.BvgError(f"bvg::vocoder_forward_cond: plan {plan_id} no longer exists")
    return plan.run_cond(x_nlc, [None if c.numel() == 0 else c for c in conds])


@_vocoder_forward_cond.register_fake
def _(x_nlc, conds, plan_id, hop):
    return x_nlc.new_empty((x_nlc.shape[0], 1, x_nlc.shape[1] * hop), dtype=torch.float32)


@torch.library.custom_op("bvg::vocoder_forward_pcm16", mutates_args=())
def _vocoder_forward_pcm16(mel: torch.Tensor, plan_id: int, hop: int) -> torch.Tensor:
    plan = _PLANS.get(plan_id)
    if plan is None:
        raise _lib.BvgError(f"bvg::vocoder_forward_pcm16: plan {plan_id} no longer exists")
    return plan.run(mel, pcm16=True)


@_vocoder_forward_pcm16.register_fake
def _(mel, plan_id, hop):
    return mel.new_empty((mel.shape[0], 1, mel.shape[2] * hop), dtype=torch.int16)


class _GeneratorPlanMixin:
    """Precision switch and the cached whole-generator launch plan, shared by the v2 ``BigVGAN`` below and the v1
    generator (``v1.py``).  Expects ``self.__dict__['_bvg_precision' | '_bvg_plan' | '_bvg_plan_key']``."""

    def _init_plan_state(self):
        self.__dict__["_bvg_precision"] = DEFAULT_PRECISION
        self.__dict__["_bvg_plan"] = None
        self.__dict__["_bvg_plan_key"] = None
        self.__dict__["_bvg_graph_max_frames"] = int(os.environ.get("BVG_GRAPH_MAX_FRAMES", "0") or 0)
        self.__dict__["_bvg_gen"] = 0
        self.__dict__["_bvg_plan_fast"] = None
        self.__dict__["_bvg_plan_calls"] = 0

    def set_cuda_graphs(self, max_frames: int):
        """Extra (not in the reference): replay calls with ``B * T_mel <= max_frames`` from CUDA graphs (one graph per
        input shape and host thread, captured on first use; 0 switches it off).  For launch-bound single-utterance
        calls such as ``infer_v2.py:860`` (B = 1); bit-identical results."""
        self.__dict__["_bvg_graph_max_frames"] = int(max_frames)
        if self.__dict__["_bvg_plan"] is not None:
            self.__dict__["_bvg_plan"].graph_max_frames = int(max_frames)
        return self

    def set_precision(self, precision: str):
        ops.precision_code(precision)
        self.__dict__["_bvg_precision"] = precision
        self._bump_generation()
        for m in self.modules():
            if isinstance(m, (Conv1d, ConvTranspose1d, AMPBlock1, AMPBlock2)):
                m.bvg_precision = precision
        return self

    @property
    def precision(self) -> str:
        return self.__dict__["_bvg_precision"]

    # ---- plan cache invalidation -------------------------------------------------------------------------------------
    # The packed plan is valid while no parameter changed.  Walking the module tree for (data_ptr, _version) of all
    # ~450-780 parameters on every forward cost 1.5-2 ms of host time -- as much as the whole single-utterance GPU step --
    # so the per-call check is reduced to the `_version` counters of a CACHED parameter list (60 us; catches every
    # in-place edit made through torch ops: `p.mul_()`, `p.copy_()`, optimizer steps), and the full walk happens only when
    # a generation counter says the parameter SET may have changed: it is bumped by every hook through which torch
    # replaces parameters or their storage (`_apply`: .to()/.cuda()/.half(); `load_state_dict`; `remove_weight_norm`;
    # `set_precision`) and by `invalidate_plan()`.  Edits through `p.data` (`p.data.copy_()`, `.data.normal_()` as in
    # `init_weights`) bump neither counter -- torch offers no hook -- so such callers call `invalidate_plan()`; as a
    # safety net the full fingerprint (incl. data pointers) is re-taken every PLAN_RECHECK_CALLS forwards.
    PLAN_RECHECK_CALLS = 64

    def invalidate_plan(self):
        """Extra (not in the reference): drop the packed-weight plan; the next forward re-folds and re-packs the weights.
        Needed only after parameter edits that bypass torch's version counters (``p.data.copy_()``,
        ``init_weights``-style ``.data.normal_()``)."""
        self._bump_generation()
        self.__dict__["_bvg_plan"] = None
        self.__dict__["_bvg_plan_key"] = None
        return self

    def _bump_generation(self):
        self.__dict__["_bvg_gen"] = self.__dict__.get("_bvg_gen", 0) + 1

    def _apply(self, fn, *args, **kwargs):
        self._bump_generation()
        return super()._apply(fn, *args, **kwargs)

    def load_state_dict(self, *args, **kwargs):
        self._bump_generation()
        return super().load_state_dict(*args, **kwargs)

    def _plan(self, device: torch.device) -> _PackedGenerator:
        d = self.__dict__
        prec = ops.precision_code(d["_bvg_precision"])
        plan = d["_bvg_plan"]
        fast = (device, prec, d.get("_bvg_gen", 0))
        if plan is not None and d.get("_bvg_plan_fast") == fast:
            d["_bvg_plan_calls"] = d.get("_bvg_plan_calls", 0) + 1
            if d["_bvg_plan_calls"] % self.PLAN_RECHECK_CALLS != 0 and \
                    [p._version for p in d["_bvg_plan_params"]] == d["_bvg_plan_versions"]:
                return plan
        params = list(self.parameters())
        key = (device, prec, tuple((p.data_ptr(), p._version) for p in params))
        if plan is None or d["_bvg_plan_key"] != key:
            for p in params:
                if p.device != device:
                    raise _lib.BvgError(f"generator parameters live on {p.device} but the input is on {device}; "
                                        "call .to(device) first")
            d["_bvg_plan"] = _PackedGenerator(self, device, prec)
            d["_bvg_plan_key"] = key
        d["_bvg_plan_fast"] = fast
        d["_bvg_plan_params"] = params
        d["_bvg_plan_versions"] = [p._version for p in params]
        return d["_bvg_plan"]


class BigVGAN(
    _GeneratorPlanMixin,
    torch.nn.Module,
    PyTorchModelHubMixin,
    library_name="bigvgan",
    repo_url="https://github.com/NVIDIA/BigVGAN",
    docs_url="https://github.com/NVIDIA/BigVGAN/blob/main/README.md",
    pipeline_tag="audio-to-audio",
    license="mit",
    tags=["neural-vocoder", "audio-generation", "arxiv:2206.04658"],
):
    """Reference bigvgan.py:243-492.

    Args:
        h (AttrDict): hyper-parameters (``config.json``); mutated like the reference: ``h["use_cuda_kernel"]``.
        use_cuda_kernel (bool): accepted and stored; both values run the sm_100a kernels (no torch path exists).

    Extra (not in the reference): ``set_precision("bf16" | "fp32" | "fp32_simt")``; default from ``$BVG_PRECISION`` or
    "bf16" (bf16 tensor-core convs with fp32 accumulation / residual stream / activation math).  "fp32" is the parity mode
    for the reference's fp32 vocoder call (infer_v2.py:830-831, 860): the same tcgen05 kernels with every operand split
    into three bf16 pieces (fp32-exact products, fp32 accumulation); "fp32_simt" is its CUDA-core cross-check.
    """

    def __init__(self, h: AttrDict, use_cuda_kernel: bool = False):
        super().__init__()
        self.h = h
        self.h["use_cuda_kernel"] = use_cuda_kernel
        Activation1d = _select_activation1d(self.h)

        self.num_kernels = len(h.resblock_kernel_sizes)
        self.num_upsamples = len(h.upsample_rates)

        self.conv_pre = weight_norm(Conv1d(h.num_mels, h.upsample_initial_channel, 7, 1, padding=3))

        if h.resblock == "1":
            resblock_class = AMPBlock1
        elif h.resblock == "2":
            resblock_class = AMPBlock2
        else:
            raise ValueError(f"Incorrect resblock class specified in hyperparameters. Got {h.resblock}")

        self.ups = nn.ModuleList()
        for i, (u, k) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
            self.ups.append(nn.ModuleList([
                weight_norm(ConvTranspose1d(h.upsample_initial_channel // (2 ** i),
                                            h.upsample_initial_channel // (2 ** (i + 1)), k, u, padding=(k - u) // 2))
            ]))

        self.resblocks = nn.ModuleList()
        for i in range(len(self.ups)):
            ch = h.upsample_initial_channel // (2 ** (i + 1))
            for j, (k, d) in enumerate(zip(h.resblock_kernel_sizes, h.resblock_dilation_sizes)):
                self.resblocks.append(resblock_class(h, ch, k, d, activation=h.activation))

        activation_post = (
            activations.Snake(ch, alpha_logscale=h.snake_logscale) if h.activation == "snake"
            else (activations.SnakeBeta(ch, alpha_logscale=h.snake_logscale) if h.activation == "snakebeta" else None))
        if activation_post is None:
            raise NotImplementedError(
                "activation incorrectly specified. check the config file and look for 'activation'.")
        self.activation_post = Activation1d(activation=activation_post)

        self.use_bias_at_final = h.get("use_bias_at_final", True)
        self.conv_post = weight_norm(Conv1d(ch, 1, 7, 1, padding=3, bias=self.use_bias_at_final))

        for i in range(len(self.ups)):
            self.ups[i].apply(init_weights)
        self.conv_post.apply(init_weights)

        self.use_tanh_at_final = h.get("use_tanh_at_final", True)
        self._init_plan_state()

    def forward(self, x):
        """``[B, num_mels, T_mel]`` CUDA tensor -> ``[B, 1, T_mel * hop]`` fp32 in [-1, 1] (reference :360-386)."""
        if x.dim() != 3 or x.shape[1] != self.h.num_mels:
            raise ValueError(f"expected mel of shape [B, {self.h.num_mels}, T], got {tuple(x.shape)}")
        if not x.is_cuda:
            raise _lib.BvgError("BigVGAN.forward: libbvg has no CPU path; move the module and the mel to a B200 "
                                f"(got device {x.device})")
        _lib.require_device(x.device.index if x.device.index is not None else torch.cuda.current_device())
        if x.dtype != torch.float32:
            x = x.float()
        plan = self._plan(x.device)
        return _vocoder_forward(x, plan.plan_id, plan.hop)

    def forward_pcm16(self, x):
        """Extra (not in the reference): ``forward`` with the caller's PCM conversion fused into the tail kernel.
        Returns the int16 ``[B, 1, T_mel * hop]`` tensor that ``infer_v2.py:860-866, 912`` compute from the waveform:
        ``torch.clamp(32767 * wav, -32767.0, 32767.0).type(torch.int16)``."""
        if x.dim() != 3 or x.shape[1] != self.h.num_mels:
            raise ValueError(f"expected mel of shape [B, {self.h.num_mels}, T], got {tuple(x.shape)}")
        if not x.is_cuda:
            raise _lib.BvgError("BigVGAN.forward_pcm16: libbvg has no CPU path; move the module and the mel to a B200 "
                                f"(got device {x.device})")
        _lib.require_device(x.device.index if x.device.index is not None else torch.cuda.current_device())
        if x.dtype != torch.float32:
            x = x.float()
        plan = self._plan(x.device)
        return _vocoder_forward_pcm16(x, plan.plan_id, plan.hop)

    def remove_weight_norm(self):
        self._bump_generation()
        try:
            print("Removing weight norm...")
            for l in self.ups:
                for l_i in l:
                    remove_weight_norm(l_i)
            for l in self.resblocks:
                l.remove_weight_norm()
            remove_weight_norm(self.conv_pre)
            remove_weight_norm(self.conv_post)
        except ValueError:
            print("[INFO] Model already removed weight norm. Skipping!")
            pass

    # ---- huggingface_hub support (reference bigvgan.py:403-492) ---------------------------------------------
    def _save_pretrained(self, save_directory: Path) -> None:
        model_path = save_directory / "bigvgan_generator.pt"
        torch.save({"generator": self.state_dict()}, model_path)
        config_path = save_directory / "config.json"
        with open(config_path, "w") as config_file:
            json.dump(self.h, config_file, indent=4)

    @classmethod
    def _from_pretrained(
            cls,
            *,
            model_id: str,
            revision: str = None,
            cache_dir: str = None,
            force_download: bool = False,
            proxies: Optional[Dict] = None,
            resume_download: bool = False,
            local_files_only: bool = False,
            token: Union[str, bool, None] = None,
            map_location: str = "cpu",
            strict: bool = False,
            use_cuda_kernel: bool = False,
            **model_kwargs,
    ):
        if os.path.isdir(model_id):
            print("Loading config.json from local directory")
            config_file = os.path.join(model_id, "config.json")
        else:
            config_file = hf_hub_download(repo_id=model_id, filename="config.json", revision=revision,
                                          cache_dir=cache_dir, force_download=force_download, proxies=proxies,
                                          resume_download=resume_download, token=token,
                                          local_files_only=local_files_only)
        h = load_hparams_from_json(config_file)
        model = cls(h, use_cuda_kernel=use_cuda_kernel)
        if os.path.isdir(model_id):
            print("Loading weights from local directory")
            model_file = os.path.join(model_id, "bigvgan_generator.pt")
        else:
            print(f"Loading weights from {model_id}")
            model_file = hf_hub_download(repo_id=model_id, filename="bigvgan_generator.pt", revision=revision,
                                         cache_dir=cache_dir, force_download=force_download, proxies=proxies,
                                         resume_download=resume_download, token=token,
                                         local_files_only=local_files_only)
        checkpoint_dict = torch.load(model_file, map_location=map_location)
        try:
            model.load_state_dict(checkpoint_dict["generator"])
        except RuntimeError:
            print("[INFO] the pretrained checkpoint does not contain weight norm. Loading the checkpoint after "
                  "removing weight norm!")
            model.remove_weight_norm()
            model.load_state_dict(checkpoint_dict["generator"])
        return model
