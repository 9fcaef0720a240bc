"""bvg_forward_noalloc (include/bvg.h): the forward plan with its scratch in the library's static device array.

No counterpart in the reference; the entry point was added for callers that cannot
call cudaMalloc.  It is checked against the existing entry point (bvg_forward with a caller workspace) through the C ABI,
on two calls in a row that share the array.  Collected last on purpose (file name), and self-contained: it passes alone in
a fresh process and after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest
import torch

from helpers import MINI_H, O, build_model, snr_db

pytestmark = pytest.mark.gpu


def _forward_workspace(lib, _lib, plan, mel):
    """Existing entry point: bvg_forward with a caller-allocated workspace."""
    B, _, T = mel.shape
    wav = torch.empty((B, 1, T * plan.hop), dtype=torch.float32, device=mel.device)
    n = lib.bvg_workspace_bytes(ctypes.byref(plan.desc), B, T)
    assert n > 0
    ws = torch.empty(n, dtype=torch.uint8, device=mel.device)
    _lib.check(lib.bvg_forward(ctypes.byref(plan.desc), mel.data_ptr(), mel.stride(0), mel.stride(1), wav.data_ptr(), B, T,
                               ws.data_ptr(), n, torch.cuda.current_stream().cuda_stream), "bvg_forward")
    torch.cuda.synchronize()
    return wav


def _forward_noalloc(lib, _lib, plan, mel, wav):
    """New entry point: inputs and outputs from the caller, no workspace argument.  `wav` is allocated by the caller
    BEFORE the call; the allocator statistics around the call show that torch was not asked for memory either."""
    B, _, T = mel.shape
    _lib.check(lib.bvg_forward_noalloc(ctypes.byref(plan.desc), mel.data_ptr(), mel.stride(0), mel.stride(1),
                                       wav.data_ptr(), B, T, torch.cuda.current_stream().cuda_stream),
               "bvg_forward_noalloc")
    torch.cuda.synchronize()
    return wav


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
def test_noalloc_forward_matches_workspace_forward_twice(precision):
    from index_tts_b200 import _lib

    lib = _lib.load()
    sd = O.make_state_dict(MINI_H, seed=5, weight_norm=False)
    m = build_model(MINI_H, sd, precision=precision)
    dev = torch.device("cuda", torch.cuda.current_device())
    plan = m._plan(dev)
    _lib.check(lib.bvg_init(-1), "bvg_init")
    g = torch.Generator().manual_seed(3)
    # two calls in a row on the shared array, the second with another (ragged, shorter) shape and other data, so that it
    # runs over whatever the first one left behind
    for call, (B, T) in enumerate([(2, 23), (1, 17)]):
        mel = (0.8 * torch.randn(B, MINI_H["num_mels"], T, generator=g)).to(dev)
        assert lib.bvg_workspace_bytes(ctypes.byref(plan.desc), B, T) <= _lib.NOALLOC_SCRATCH_BYTES
        ref = _forward_workspace(lib, _lib, plan, mel)
        wav = torch.full_like(ref, float("nan"))
        before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
        _forward_noalloc(lib, _lib, plan, mel, wav)
        assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before
        a, b = ref.cpu().numpy(), wav.cpu().numpy()
        assert np.isfinite(b).all()
        err = float(np.abs(a - b).max())
        print(f"noalloc {precision} call {call} B={B} T={T}: max-abs {err:.3e} vs bvg_forward, "
              f"bit-identical={np.array_equal(a, b)}")
        # same kernels, same inputs: the tolerance of tests/test_gpu_forward.py is the bar, identity is the expectation
        assert err <= 1e-4
        assert snr_db(a, b) >= 80.0


def test_noalloc_full_generator_and_size_limit():
    """The 22 kHz generator (every kernel variant of the plan, tensor-core Activation1d included) and the size check."""
    from index_tts_b200 import _lib

    lib = _lib.load()
    sd = O.make_state_dict(O.DEFAULT_HPARAMS, seed=11)
    m = build_model(O.DEFAULT_HPARAMS, sd, precision="bf16")
    dev = torch.device("cuda", torch.cuda.current_device())
    plan = m._plan(dev)
    g = torch.Generator().manual_seed(4)
    for B, T in [(2, 40), (1, 64)]:
        mel = (0.8 * torch.randn(B, O.DEFAULT_HPARAMS["num_mels"], T, generator=g)).to(dev)
        ref = _forward_workspace(lib, _lib, plan, mel)
        wav = _forward_noalloc(lib, _lib, plan, mel, torch.full_like(ref, float("nan")))
        a, b = ref.cpu().numpy(), wav.cpu().numpy()
        err = float(np.abs(a - b).max())
        print(f"noalloc full generator B={B} T={T}: max-abs {err:.3e}, bit-identical={np.array_equal(a, b)}")
        assert np.isfinite(b).all() and err <= 1e-4 and snr_db(a, b) >= 80.0

    # beyond the built-in array: a clear error, nothing launched, output untouched
    B, T = 4, 2048
    assert lib.bvg_workspace_bytes(ctypes.byref(plan.desc), B, T) > _lib.NOALLOC_SCRATCH_BYTES
    mel = torch.zeros(B, O.DEFAULT_HPARAMS["num_mels"], T, device=dev)
    wav = torch.full((B, 1, T * plan.hop), 7.0, device=dev)
    rc = lib.bvg_forward_noalloc(ctypes.byref(plan.desc), mel.data_ptr(), mel.stride(0), mel.stride(1), wav.data_ptr(), B, T,
                                 torch.cuda.current_stream().cuda_stream)
    assert rc == -5  # BVG_ERR_WORKSPACE
    msg = lib.bvg_last_error().decode()
    assert "bvg_forward_noalloc" in msg and str(_lib.NOALLOC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert bool((wav == 7.0).all())
