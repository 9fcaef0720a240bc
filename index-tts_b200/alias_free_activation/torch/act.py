"""``Activation1d`` with the reference's torch-path constructor (alias_free_activation/torch/act.py:8-30).

The reference has two implementations (unfused torch ops / fused CUDA extension) selected by ``use_cuda_kernel``;
here both names resolve to the one fused sm_100a kernel -- there is no torch path to fall back to."""
from .._fused import FusedActivation1dBase


class Activation1d(FusedActivation1dBase):
    def __init__(self, activation, up_ratio: int = 2, down_ratio: int = 2, up_kernel_size: int = 12,
                 down_kernel_size: int = 12):
        super().__init__(activation, up_ratio, down_ratio, up_kernel_size, down_kernel_size)


__all__ = ["Activation1d"]
