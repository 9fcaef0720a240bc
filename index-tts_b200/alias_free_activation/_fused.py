"""Shared implementation of ``Activation1d`` for both import paths the reference exposes
(``alias_free_activation.torch.act`` and ``alias_free_activation.cuda.activation1d``): one fused sm_100a kernel."""
import torch
import torch.nn as nn

from .. import ops


class FusedActivation1dBase(nn.Module):
    def __init__(self, activation, up_ratio: int = 2, down_ratio: int = 2, up_kernel_size: int = 12,
                 down_kernel_size: int = 12):
        super().__init__()
        from .torch.resample import DownSample1d, UpSample1d  # late import: avoids an import cycle

        if (up_ratio, down_ratio, up_kernel_size, down_kernel_size) != (2, 2, 12, 12):
            # the reference kernel silently computes the wrong thing here (cuda/activation1d.py:17-19 NOTE)
            raise NotImplementedError("the fused Activation1d kernel implements ratio 2 / kernel size 12 only")
        self.up_ratio = up_ratio
        self.down_ratio = down_ratio
        self.act = activation
        self.upsample = UpSample1d(up_ratio, up_kernel_size)
        self.downsample = DownSample1d(down_ratio, down_kernel_size)
        self.precise = None  # None: accurate sin for fp32 tensors, MUFU approximations for bf16

    def log_params(self):
        """alpha/beta in log scale as the kernel expects them (reference cuda/activation1d.py:61-72)."""
        alpha = self.act.alpha.data
        beta = alpha if self.act.__class__.__name__ == "Snake" else self.act.beta.data
        if not self.act.alpha_logscale:
            alpha, beta = torch.log(alpha), torch.log(beta)
        return alpha, beta

    # x: [B, C, T]
    def forward(self, x):
        alpha, beta = self.log_params()
        precise = (x.dtype == torch.float32) if self.precise is None else bool(self.precise)
        return ops.activation1d(x, alpha, beta, precise)
