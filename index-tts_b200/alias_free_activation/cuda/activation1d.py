"""Fused ``Activation1d`` (reference alias_free_activation/cuda/activation1d.py:14-77).

``infer_v2.py:174-182`` imports this module and reads ``activation1d.anti_alias_activation_cuda`` to decide whether
the fused kernel is usable; importing it here loads ``libbvg.so`` and raises if that is impossible (the reference
swallows the failure and silently falls back to torch)."""
import torch

from . import load
from .._fused import FusedActivation1dBase

anti_alias_activation_cuda = load.load()


class FusedAntiAliasActivation(torch.autograd.Function):
    """Same contract as the reference (activation1d.py:14-32): filter size 12, replicate padding, log-scale
    alpha/beta; backward is not implemented."""

    @staticmethod
    def forward(ctx, inputs, up_ftr, down_ftr, alpha, beta):
        return anti_alias_activation_cuda.forward(inputs, up_ftr, down_ftr, alpha, beta)

    @staticmethod
    def backward(ctx, output_grads):
        raise NotImplementedError


class Activation1d(FusedActivation1dBase):
    def __init__(self, activation, up_ratio: int = 2, down_ratio: int = 2, up_kernel_size: int = 12,
                 down_kernel_size: int = 12, fused: bool = True):
        super().__init__(activation, up_ratio, down_ratio, up_kernel_size, down_kernel_size)
        self.fused = fused  # kept for API parity; fused=False has no separate implementation here
