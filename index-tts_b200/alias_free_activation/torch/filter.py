"""Filter design and the ``LowPassFilter1d`` buffer holder (reference alias_free_activation/torch/filter.py).

``kaiser_sinc_filter1d`` is host-side filter *design* (12 numbers); the filtering itself happens inside the fused
kernel, which bakes the taps of the (cutoff 0.25, half-width 0.3, 12-tap) design as constants.  The ``filter``
buffers are registered so state dicts keep the reference's 783/667 keys.
"""
import math

import torch
import torch.nn as nn


def kaiser_sinc_filter1d(cutoff, half_width, kernel_size):
    """Reference filter.py:30-62; returns ``[1, 1, kernel_size]``."""
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
    window = torch.kaiser_window(kernel_size, beta=beta, periodic=False)
    if even:
        time = torch.arange(-half_size, half_size) + 0.5
    else:
        time = torch.arange(kernel_size) - half_size
    if cutoff == 0:
        filter_ = torch.zeros_like(time)
    else:
        filter_ = 2 * cutoff * window * torch.sinc(2 * cutoff * time)
        filter_ /= filter_.sum()
    return filter_.view(1, 1, kernel_size)


class LowPassFilter1d(nn.Module):
    """Buffer holder with the reference's constructor (filter.py:65-92).  Not callable on its own: the low-pass is
    fused into ``Activation1d``'s kernel."""

    def __init__(self, cutoff=0.5, half_width=0.6, stride: int = 1, padding: bool = True,
                 padding_mode: str = "replicate", kernel_size: int = 12):
        super().__init__()
        if cutoff < -0.0:
            raise ValueError("Minimum cutoff must be larger than zero.")
        if cutoff > 0.5:
            raise ValueError("A cutoff above 0.5 does not make sense.")
        self.kernel_size = kernel_size
        self.even = kernel_size % 2 == 0
        self.pad_left = kernel_size // 2 - int(self.even)
        self.pad_right = kernel_size // 2
        self.stride = stride
        self.padding = padding
        self.padding_mode = padding_mode
        self.register_buffer("filter", kaiser_sinc_filter1d(cutoff, half_width, kernel_size))

    def forward(self, x):
        raise NotImplementedError("LowPassFilter1d is fused into Activation1d's sm_100a kernel; call Activation1d")
