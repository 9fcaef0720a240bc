"""Replacement for reference alias_free_activation/cuda/load.py:17-65 (a 6-minute ninja JIT at first import): the
library is prebuilt in-tree (``index-tts_b200/libbvg.so``) and loaded with ctypes.  ``load()`` returns an object
with the reference extension's single entry point ``forward(inputs, up_ftr, down_ftr, alpha, beta)``."""
import torch

from ... import _lib, ops


class _AntiAliasActivationCuda:
    """Stand-in for the pybind module ``anti_alias_activation_cuda`` (anti_alias_activation.cpp:19-23)."""

    def __init__(self):
        self.lib = _lib.load()
        self.path = _lib.LIB_PATH

    def forward(self, inputs: torch.Tensor, up_ftr, down_ftr, alpha: torch.Tensor, beta: torch.Tensor) -> torch.Tensor:
        # the filters are compile-time constants of the kernel; the arguments are accepted for call compatibility
        precise = inputs.dtype == torch.float32
        return ops.activation1d(inputs, alpha, beta, precise)

    def __repr__(self):
        return f"<libbvg anti_alias_activation_cuda at {self.path}>"


def load():
    return _AntiAliasActivationCuda()
