"""ctypes binding of ``libbvg.so`` (the C ABI declared in ``include/bvg.h``).

There is no fallback: if the shared library is missing or the device is not sm_100, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbvg.so")
if os.environ.get("BVG_LIB_PATH"):  # developer A/B of kernel build variants (tools/build_variant.sh)
    LIB_PATH = os.environ["BVG_LIB_PATH"]

BVG_F32, BVG_BF16, BVG_I16, BVG_BF16X3 = 0, 1, 2, 3
NOALLOC_SCRATCH_BYTES = 64 << 20  # BVG_NOALLOC_SCRATCH_BYTES (include/bvg.h): scratch limit of bvg_forward_noalloc
BVG_EXACT_F32, BVG_BF16_TC, BVG_F32_SPLIT_TC = 0, 1, 2
SPLIT_TERMS = 6  # term slots of a BVG_BF16X3 tensor (include/bvg.h)
MAX_STAGES, MAX_KERNELS, MAX_DILATIONS, MAX_PHASES = 8, 4, 4, 8

c_i64, c_int, c_f32, c_vp, c_sz = C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_size_t


class ConvDesc(C.Structure):
    _fields_ = [("w", c_vp), ("bias", c_vp), ("cin", C.c_int32), ("cout", C.c_int32), ("k", C.c_int32),
                ("dilation", C.c_int32), ("fold", C.c_int32), ("reserved", C.c_int32)]


class ActDesc(C.Structure):
    _fields_ = [("alpha_log", c_vp), ("beta_log", c_vp)]


class ModelDesc(C.Structure):
    _fields_ = [
        ("num_mels", C.c_int32),
        ("upsample_initial_channel", C.c_int32),
        ("num_stages", C.c_int32),
        ("num_kernels", C.c_int32),
        ("num_dilations", C.c_int32),
        ("upsample_rates", C.c_int32 * MAX_STAGES),
        ("upsample_kernel_sizes", C.c_int32 * MAX_STAGES),
        ("use_tanh_at_final", C.c_int32),
        ("precision", C.c_int32),
        ("resblock", C.c_int32),
        ("conv_pre", ConvDesc),
        ("ups", ConvDesc * MAX_STAGES),
        ("convs1", ((ConvDesc * MAX_DILATIONS) * MAX_KERNELS) * MAX_STAGES),
        ("convs2", ((ConvDesc * MAX_DILATIONS) * MAX_KERNELS) * MAX_STAGES),
        ("acts", ((ActDesc * (2 * MAX_DILATIONS)) * MAX_KERNELS) * MAX_STAGES),
        ("act_post", ActDesc),
        ("conv_post", ConvDesc),
    ]


class LaunchRecord(C.Structure):
    _fields_ = [("kind", C.c_int32), ("stage", C.c_int32), ("cin", C.c_int32), ("cout", C.c_int32), ("k", C.c_int32),
                ("dilation", C.c_int32), ("rows", C.c_int64), ("flops", C.c_double), ("bytes", C.c_double),
                ("ms", C.c_float)]


KIND_NAMES = {0: "layout", 1: "conv1d", 2: "convtranspose1d", 3: "activation1d", 4: "conv_post", 5: "cond_add"}

# name -> (restype, argtypes); must list every symbol include/bvg.h declares (tests/test_abi.py checks that)
SIGNATURES = {
    "bvg_last_error": (C.c_char_p, []),
    "bvg_version": (c_int, []),
    "bvg_device_check": (c_int, [c_int]),
    "bvg_launch_count": (c_i64, []),
    "bvg_reset_launch_count": (None, []),
    "bvg_init": (c_int, [c_int]),
    "bvg_activation1d_fwd": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_i64, c_int, c_int, c_vp]),
    "bvg_activation1d_nlc_fwd": (c_int, [c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_vp]),
    "bvg_activation1d_nlc_multi_fwd": (c_int, [c_vp, c_int, C.POINTER(c_vp), c_int, C.POINTER(c_vp), C.POINTER(c_vp), c_int,
                                               c_i64, c_i64, c_i64, c_int, c_vp]),
    "bvg_fold_weight_norm": (c_int, [c_vp, c_vp, c_vp, c_i64, c_i64, c_vp]),
    "bvg_pack_conv1d_weight": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_vp]),
    "bvg_pack_convtranspose1d_weight": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_vp]),
    "bvg_split_bf16x3_nlc": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp]),
    "bvg_ncl_to_nlc": (c_int, [c_vp, c_int, c_i64, c_i64, c_vp, c_int, c_i64, c_i64, c_i64, c_vp]),
    "bvg_nlc_to_ncl": (c_int, [c_vp, c_int, c_vp, c_int, c_i64, c_i64, c_i64, c_vp]),
    "bvg_conv1d_nlc_fwd": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_int,
                                   c_f32, c_int, c_vp]),
    "bvg_convtranspose1d_nlc_fwd": (c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_int,
                                            c_int, c_vp]),
    "bvg_conv_post_fwd": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_int, c_vp]),
    "bvg_conv_post_pcm16_fwd": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_int, c_int, c_vp]),
    "bvg_forward_pcm16": (c_int, [C.POINTER(ModelDesc), c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_sz, c_vp]),
    "bvg_workspace_bytes": (c_sz, [C.POINTER(ModelDesc), c_i64, c_i64]),
    "bvg_forward": (c_int, [C.POINTER(ModelDesc), c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_sz, c_vp]),
    "bvg_forward_noalloc": (c_int, [C.POINTER(ModelDesc), c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp]),
    "bvg_forward_profile": (c_int, [C.POINTER(ModelDesc), c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_sz, c_vp,
                                    C.POINTER(LaunchRecord), c_int, C.POINTER(c_int)]),
    "bvg_forward_cond_profile": (c_int, [C.POINTER(ModelDesc), c_vp, c_int, c_i64, c_i64, C.POINTER(c_vp), c_vp, c_i64, c_i64,
                                         c_vp, c_sz, c_vp, C.POINTER(LaunchRecord), c_int, C.POINTER(c_int)]),
    "bvg_debug_conv_trace": (c_int, [c_vp, c_int]),
    "bvg_cond_linear_fwd": (c_int, [c_vp, c_vp, c_vp, c_vp, c_i64, c_i64, c_i64, c_vp]),
    "bvg_add_channel_vector_nlc": (c_int, [c_vp, c_int, c_vp, c_vp, c_int, c_i64, c_i64, c_i64, c_vp]),
    "bvg_forward_cond": (c_int, [C.POINTER(ModelDesc), c_vp, c_int, c_i64, c_i64, C.POINTER(c_vp), c_vp, c_i64, c_i64,
                                 c_vp, c_sz, c_vp]),
}

_lock = threading.Lock()
_lib = None


class BvgError(RuntimeError):
    pass


def load():
    """Load ``libbvg.so`` (once).  Raises ``BvgError`` if it has not been built -- there is no Python/torch fallback."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise BvgError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or index-tts_b200/csrc/build.sh). This package has no CPU or torch fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str = "bvg"):
    if rc != 0:
        msg = load().bvg_last_error()
        raise BvgError(f"{what} failed with status {rc}: {msg.decode() if msg else ''}")


_device_ok = set()


def require_device(index: int):
    """Hard error unless CUDA device ``index`` is sm_100 (replaces the reference's silent fallback,
    infer_v2.py:176-182)."""
    if index in _device_ok:
        return
    check(load().bvg_device_check(int(index)), "bvg_device_check")
    _device_ok.add(index)
