"""Torch-facing wrappers over the C ABI (``include/bvg.h``).

Torch is plumbing here: it owns device memory and the current stream; every computation is a call into
``libbvg.so`` through ctypes.  The ops are also registered with ``torch.library`` (namespace ``bvg``) with fake
(meta) implementations so that ``torch.compile(dynamic=True)`` (reference infer_v2.py:267-270) can trace through
the module.  There is no autograd formula (the reference extension raises in backward, activation1d.py:29-32)
and no CPU implementation: a non-CUDA tensor is an error.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib
from ._lib import (BVG_BF16, BVG_BF16_TC, BVG_BF16X3, BVG_EXACT_F32, BVG_F32, BVG_F32_SPLIT_TC, SPLIT_TERMS, check)

BF16X3 = "bf16x3"  # out_dtype sentinel: the split operand tensor of the fp32 parity mode (a bf16 tensor [.., 6*C])

_DT = {torch.float32: BVG_F32, torch.bfloat16: BVG_BF16}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _require_cuda(t: torch.Tensor, what: str):
    if not t.is_cuda:
        raise _lib.BvgError(f"{what}: expected a CUDA tensor, got device {t.device} (libbvg has no CPU path)")
    _lib.require_device(t.device.index if t.device.index is not None else torch.cuda.current_device())


def _dt(t: torch.Tensor, what: str) -> int:
    if t.dtype not in _DT:
        raise _lib.BvgError(f"{what}: dtype {t.dtype} not supported (float32 and bfloat16 only)")
    return _DT[t.dtype]


def precision_code(name) -> int:
    """'bf16': tensor-core throughput mode.  'fp32': the fp32 parity mode, on the tensor cores with bf16x3 split operands
    (BVG_F32_SPLIT_TC).  'fp32_simt': the CUDA-core cross-check implementation of the fp32 mode (BVG_EXACT_F32)."""
    if isinstance(name, str):
        if name in ("fp32", "float32", "exact"):
            return BVG_F32_SPLIT_TC
        if name in ("fp32_simt", "exact_simt"):
            return BVG_EXACT_F32
        if name in ("bf16", "bfloat16"):
            return BVG_BF16_TC
    elif name is torch.float32:
        return BVG_F32_SPLIT_TC
    elif name is torch.bfloat16:
        return BVG_BF16_TC
    elif name in (BVG_EXACT_F32, BVG_BF16_TC, BVG_F32_SPLIT_TC):
        return int(name)
    raise ValueError(f"unknown precision {name!r} (use 'bf16', 'fp32' or 'fp32_simt')")


def operand_dtype(precision: int):
    """What a conv of this precision reads: torch.bfloat16, the BF16X3 sentinel or torch.float32."""
    return torch.bfloat16 if precision == BVG_BF16_TC else (BF16X3 if precision == BVG_F32_SPLIT_TC else torch.float32)


def split_bf16x3(x: torch.Tensor) -> torch.Tensor:
    """fp32 ``[..., C]`` -> the BVG_BF16X3 operand tensor ``[..., 6*C]`` (bf16; include/bvg.h)."""
    _require_cuda(x, "split_bf16x3")
    x = x.float().contiguous()
    C = x.shape[-1]
    y = torch.empty(tuple(x.shape[:-1]) + (SPLIT_TERMS * C,), dtype=torch.bfloat16, device=x.device)
    rows = x.numel() // max(C, 1)
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_split_bf16x3_nlc(x.data_ptr(), y.data_ptr(), rows, C, _stream()), "bvg_split_bf16x3_nlc")
    return y


# ---------------------------------------------------------------------------------------------------------------
# Activation1d
# ---------------------------------------------------------------------------------------------------------------
def _activation1d_impl(x: torch.Tensor, alpha_log: torch.Tensor, beta_log: torch.Tensor, precise: bool) -> torch.Tensor:
    _require_cuda(x, "activation1d")
    if x.dim() != 3:
        raise _lib.BvgError(f"activation1d: expected [B, C, T], got shape {tuple(x.shape)}")
    dt = _dt(x, "activation1d")
    B, C, T = x.shape
    if T > 0 and x.stride(2) != 1:
        x = x.contiguous()
    y = torch.empty((B, C, T), dtype=x.dtype, device=x.device)
    if B * C * T == 0:
        return y
    a = alpha_log.detach().to(device=x.device, dtype=torch.float32).contiguous()
    b = beta_log.detach().to(device=x.device, dtype=torch.float32).contiguous()
    if a.numel() != C or b.numel() != C:
        raise _lib.BvgError(f"activation1d: alpha/beta must have {C} elements")
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_activation1d_fwd(x.data_ptr(), y.data_ptr(), a.data_ptr(), b.data_ptr(), B, C, T,
                                               x.stride(0), x.stride(1), dt, int(bool(precise)), _stream()),
              "bvg_activation1d_fwd")
    return y


@torch.library.custom_op("bvg::activation1d", mutates_args=())
def activation1d(x: torch.Tensor, alpha_log: torch.Tensor, beta_log: torch.Tensor, precise: bool) -> torch.Tensor:
    """Fused up2 -> SnakeBeta -> down2 on ``[B, C, T]`` (same call shape as the reference's
    ``anti_alias_activation_cuda.forward`` minus the constant filter arguments)."""
    return _activation1d_impl(x, alpha_log, beta_log, precise)


@activation1d.register_fake
def _(x, alpha_log, beta_log, precise):
    return torch.empty_like(x, memory_format=torch.contiguous_format)


def activation1d_nlc(x: torch.Tensor, alpha_log: torch.Tensor, beta_log: torch.Tensor, out_dtype,
                     precise: bool) -> torch.Tensor:
    """Channels-last ``[B, T, C]`` variant used inside the generator.  ``out_dtype`` may be ``ops.BF16X3`` (fp32 input
    only): the result is then the split operand tensor ``[B, T, 6*C]`` of the fp32 parity mode."""
    _require_cuda(x, "activation1d_nlc")
    x = x.contiguous()
    B, T, C = x.shape
    if out_dtype == BF16X3:
        y = torch.empty((B, T, SPLIT_TERMS * C), dtype=torch.bfloat16, device=x.device)
        y_code = BVG_BF16X3
    else:
        y = torch.empty((B, T, C), dtype=out_dtype, device=x.device)
        y_code = _DT[out_dtype]
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_activation1d_nlc_fwd(x.data_ptr(), _dt(x, "activation1d_nlc"), y.data_ptr(), y_code,
                                                   alpha_log.data_ptr(), beta_log.data_ptr(), B, T, C,
                                                   int(bool(precise)), _stream()), "bvg_activation1d_nlc_fwd")
    return y


def activation1d_nlc_multi(x: torch.Tensor, alphas_log, betas_log, out_dtype: torch.dtype, precise: bool):
    """``[Activation1d(x; alphas_log[h], betas_log[h]) for h]`` on a channels-last ``[B, T, C]`` tensor: the first
    activations of the AMP blocks of one stage (bigvgan.py:369-374, :133-135) share their up-sampler in one launch."""
    import ctypes
    _require_cuda(x, "activation1d_nlc_multi")
    x = x.contiguous()
    B, T, C = x.shape
    n = len(alphas_log)
    if n == 0 or len(betas_log) != n:
        raise _lib.BvgError("activation1d_nlc_multi: need as many beta as alpha tensors (at least one)")
    ys = [torch.empty((B, T, C), dtype=out_dtype, device=x.device) for _ in range(n)]
    arr = ctypes.c_void_p * n
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_activation1d_nlc_multi_fwd(
            x.data_ptr(), _dt(x, "activation1d_nlc_multi"), arr(*[y.data_ptr() for y in ys]), _DT[out_dtype],
            arr(*[a.data_ptr() for a in alphas_log]), arr(*[b.data_ptr() for b in betas_log]), n, B, T, C,
            int(bool(precise)), _stream()), "bvg_activation1d_nlc_multi_fwd")
    return ys


# ---------------------------------------------------------------------------------------------------------------
# layout + weights
# ---------------------------------------------------------------------------------------------------------------
def ncl_to_nlc(x: torch.Tensor, out_dtype: torch.dtype) -> torch.Tensor:
    _require_cuda(x, "ncl_to_nlc")
    if x.stride(2) != 1:
        x = x.contiguous()
    B, Cc, T = x.shape
    y = torch.empty((B, T, Cc), dtype=out_dtype, device=x.device)
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_ncl_to_nlc(x.data_ptr(), _dt(x, "ncl_to_nlc"), x.stride(0), x.stride(1), y.data_ptr(),
                                         _DT[out_dtype], B, Cc, T, _stream()), "bvg_ncl_to_nlc")
    return y


def nlc_to_ncl(x: torch.Tensor, out_dtype: torch.dtype) -> torch.Tensor:
    _require_cuda(x, "nlc_to_ncl")
    x = x.contiguous()
    B, T, Cc = x.shape
    y = torch.empty((B, Cc, T), dtype=out_dtype, device=x.device)
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_nlc_to_ncl(x.data_ptr(), _dt(x, "nlc_to_ncl"), y.data_ptr(), _DT[out_dtype], B, Cc, T,
                                         _stream()), "bvg_nlc_to_ncl")
    return y


def fold_weight_norm(v: torch.Tensor, g: torch.Tensor) -> torch.Tensor:
    """``g * v / ||v||`` over all dims but 0 (reference bigvgan.py:388-400 via torch's remove_weight_norm)."""
    _require_cuda(v, "fold_weight_norm")
    v = v.detach().float().contiguous()
    g = g.detach().float().contiguous()
    w = torch.empty_like(v)
    rows = v.shape[0]
    with torch.cuda.device(v.device):
        check(_lib.load().bvg_fold_weight_norm(v.data_ptr(), g.data_ptr(), w.data_ptr(), rows, v.numel() // max(rows, 1),
                                               _stream()), "bvg_fold_weight_norm")
    return w


def pack_conv1d_weight(w: torch.Tensor, precision: int) -> torch.Tensor:
    """``[Cout, Cin, k]`` fp32 -> ``[k, Cout, Cin]`` in the operand dtype of ``precision``."""
    _require_cuda(w, "pack_conv1d_weight")
    w = w.detach().float().contiguous()
    Cout, Cin, k = w.shape
    if precision == BVG_F32_SPLIT_TC:
        out = torch.empty((k, Cout, SPLIT_TERMS * Cin), dtype=torch.bfloat16, device=w.device)
        code = BVG_BF16X3
    else:
        dt = torch.bfloat16 if precision == BVG_BF16_TC else torch.float32
        out = torch.empty((k, Cout, Cin), dtype=dt, device=w.device)
        code = _DT[dt]
    with torch.cuda.device(w.device):
        check(_lib.load().bvg_pack_conv1d_weight(w.data_ptr(), out.data_ptr(), Cout, Cin, k, code, _stream()),
              "bvg_pack_conv1d_weight")
    return out


def pack_convtranspose1d_weight(w: torch.Tensor, stride: int, precision: int) -> torch.Tensor:
    """``[Cin, Cout, K]`` fp32 -> polyphase ``[K/u, u*Cout, Cin]``."""
    _require_cuda(w, "pack_convtranspose1d_weight")
    w = w.detach().float().contiguous()
    Cin, Cout, K = w.shape
    if precision == BVG_F32_SPLIT_TC:
        out = torch.empty((K // stride, stride * Cout, SPLIT_TERMS * Cin), dtype=torch.bfloat16, device=w.device)
        code = BVG_BF16X3
    else:
        dt = torch.bfloat16 if precision == BVG_BF16_TC else torch.float32
        out = torch.empty((K // stride, stride * Cout, Cin), dtype=dt, device=w.device)
        code = _DT[dt]
    with torch.cuda.device(w.device):
        check(_lib.load().bvg_pack_convtranspose1d_weight(w.data_ptr(), out.data_ptr(), Cin, Cout, K, stride, code,
                                                          _stream()), "bvg_pack_convtranspose1d_weight")
    return out


# ---------------------------------------------------------------------------------------------------------------
# convolutions (channels-last)
# ---------------------------------------------------------------------------------------------------------------
def conv1d_nlc(x: torch.Tensor, w_packed: torch.Tensor, bias: Optional[torch.Tensor], dilation: int, precision: int,
               residual: Optional[torch.Tensor] = None, accum: Optional[torch.Tensor] = None, out_scale: float = 1.0,
               out_f32: bool = True, out_bf16: bool = False):
    _require_cuda(x, "conv1d_nlc")
    x = x.contiguous()
    B, T, Cin = x.shape
    k, Cout, Cin2 = w_packed.shape
    if Cin2 != Cin:
        raise _lib.BvgError(f"conv1d_nlc: weight has {Cin2} input channels, x has {Cin}")
    want = torch.float32 if precision == BVG_EXACT_F32 else torch.bfloat16
    if x.dtype != want or w_packed.dtype != want:
        raise _lib.BvgError(f"conv1d_nlc: operands must be {want} for this precision (x {x.dtype}, w {w_packed.dtype})")
    if precision == BVG_F32_SPLIT_TC:  # operands are BVG_BF16X3 tensors: the C ABI takes the logical channel count
        if Cin % SPLIT_TERMS or out_bf16:
            raise _lib.BvgError("conv1d_nlc: split-operand mode needs [B, T, 6*Cin] operands and has fp32 outputs only")
        Cin //= SPLIT_TERMS
    y32 = torch.empty((B, T, Cout), dtype=torch.float32, device=x.device) if out_f32 else None
    y16 = torch.empty((B, T, Cout), dtype=torch.bfloat16, device=x.device) if out_bf16 else None
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_conv1d_nlc_fwd(x.data_ptr(), w_packed.data_ptr(), _ptr(bias), _ptr(residual), _ptr(accum),
                                             _ptr(y32), _ptr(y16), B, T, Cin, Cout, k, dilation, float(out_scale),
                                             precision, _stream()), "bvg_conv1d_nlc_fwd")
    if out_f32 and out_bf16:
        return y32, y16
    return y32 if out_f32 else y16


def convtranspose1d_nlc(x: torch.Tensor, w_packed: torch.Tensor, bias: Optional[torch.Tensor], K: int, stride: int,
                        precision: int, out_f32: bool = True, out_bf16: bool = False):
    _require_cuda(x, "convtranspose1d_nlc")
    x = x.contiguous()
    B, Tin, Cin = x.shape
    taps, N, Cin2 = w_packed.shape
    Cout = N // stride
    if Cin2 != Cin or taps * stride != K:
        raise _lib.BvgError("convtranspose1d_nlc: packed weight does not match x / K / stride")
    if precision == BVG_F32_SPLIT_TC:
        if Cin % SPLIT_TERMS or out_bf16:
            raise _lib.BvgError("convtranspose1d_nlc: split-operand mode needs [B, T, 6*Cin] operands, fp32 outputs only")
        Cin //= SPLIT_TERMS
    y32 = torch.empty((B, Tin * stride, Cout), dtype=torch.float32, device=x.device) if out_f32 else None
    y16 = torch.empty((B, Tin * stride, Cout), dtype=torch.bfloat16, device=x.device) if out_bf16 else None
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_convtranspose1d_nlc_fwd(x.data_ptr(), w_packed.data_ptr(), _ptr(bias), _ptr(y32), _ptr(y16),
                                                      B, Tin, Cin, Cout, K, stride, precision, _stream()),
              "bvg_convtranspose1d_nlc_fwd")
    if out_f32 and out_bf16:
        return y32, y16
    return y32 if out_f32 else y16


def conv_post(x_nlc: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], use_tanh: bool,
              pcm16: bool = False) -> torch.Tensor:
    """``conv_post`` + clamp / tanh (reference bigvgan.py:379-384) on a channels-last fp32 tensor.  ``pcm16=True`` also
    applies the caller's ``clamp(32767 * wav, -32767, 32767).type(torch.int16)`` (infer_v2.py:866, 912) in the kernel."""
    _require_cuda(x_nlc, "conv_post")
    x_nlc = x_nlc.contiguous()
    B, T, Cin = x_nlc.shape
    w = w.detach().float().contiguous()
    y = torch.empty((B, 1, T), dtype=torch.int16 if pcm16 else torch.float32, device=x_nlc.device)
    fn = _lib.load().bvg_conv_post_pcm16_fwd if pcm16 else _lib.load().bvg_conv_post_fwd
    with torch.cuda.device(x_nlc.device):
        check(fn(x_nlc.data_ptr(), w.data_ptr(), _ptr(bias), y.data_ptr(), B, T, Cin, w.shape[-1], int(bool(use_tanh)),
                 _stream()), "bvg_conv_post_pcm16_fwd" if pcm16 else "bvg_conv_post_fwd")
    return y


# ---------------------------------------------------------------------------------------------------------------
# v1 generator conditioning (reference indextts/BigVGAN/models.py:197-202, 236, 244-245)
# ---------------------------------------------------------------------------------------------------------------
def cond_linear(e: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """``Conv1d(D, C, 1)`` applied to the length-1 speaker embedding: ``e`` ``[B, D]`` fp32, ``w`` ``[C, D]`` or
    ``[C, D, 1]`` -> ``[B, C]`` fp32."""
    _require_cuda(e, "cond_linear")
    e = e.detach().float().contiguous()
    w = w.detach().float().contiguous()
    B, D = e.shape
    C = w.shape[0]
    if w.numel() != C * D:
        raise _lib.BvgError(f"cond_linear: weight {tuple(w.shape)} does not match an embedding of width {D}")
    b = None if bias is None else bias.detach().float().contiguous()
    y = torch.empty((B, C), dtype=torch.float32, device=e.device)
    with torch.cuda.device(e.device):
        check(_lib.load().bvg_cond_linear_fwd(e.data_ptr(), w.data_ptr(), _ptr(b), y.data_ptr(), B, D, C, _stream()),
              "bvg_cond_linear_fwd")
    return y


def add_channel_vector_nlc(x: torch.Tensor, v: Optional[torch.Tensor], out_dtype: torch.dtype) -> torch.Tensor:
    """``x[b, t, c] + v[b, c]`` on a channels-last tensor (``v`` None: dtype conversion only)."""
    _require_cuda(x, "add_channel_vector_nlc")
    x = x.contiguous()
    B, T, C = x.shape
    if v is not None:
        v = v.detach().float().contiguous()
        if tuple(v.shape) != (B, C):
            raise _lib.BvgError(f"add_channel_vector_nlc: vector must be [{B}, {C}], got {tuple(v.shape)}")
    y = torch.empty((B, T, C), dtype=out_dtype, device=x.device)
    with torch.cuda.device(x.device):
        check(_lib.load().bvg_add_channel_vector_nlc(x.data_ptr(), _dt(x, "add_channel_vector_nlc"), _ptr(v), y.data_ptr(),
                                                     _DT[out_dtype], B, T, C, _stream()), "bvg_add_channel_vector_nlc")
    return y


def launch_count() -> int:
    return int(_lib.load().bvg_launch_count())


def reset_launch_count():
    _lib.load().bvg_reset_launch_count()
