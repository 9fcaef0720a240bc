"""Parameter containers for Snake / SnakeBeta (reference ``bigvgan/activations.py:9-120``).

These modules only *store* ``alpha`` / ``beta``: ``Activation1d`` hands them to the fused sm_100a kernel.  The reference
never applies a bare ``Snake`` / ``SnakeBeta`` (always through ``Activation1d``: bigvgan.py:94-130, 343-357), and this
package has no torch compute path, so calling ``forward`` on the bare module raises.
"""
import torch
from torch import nn
from torch.nn import Parameter


class Snake(nn.Module):
    """x + 1/a * sin^2(a x), per channel (reference activations.py:9-59)."""

    def __init__(self, in_features, alpha=1.0, alpha_trainable=True, alpha_logscale=False):
        super().__init__()
        self.in_features = in_features
        self.alpha_logscale = alpha_logscale
        if self.alpha_logscale:
            self.alpha = Parameter(torch.zeros(in_features) * alpha)
        else:
            self.alpha = Parameter(torch.ones(in_features) * alpha)
        self.alpha.requires_grad = alpha_trainable
        self.no_div_by_zero = 0.000000001

    def forward(self, x):
        raise NotImplementedError("Snake is a parameter holder here: wrap it in Activation1d (the fused sm_100a kernel); "
                                  "this package has no torch compute path")


class SnakeBeta(nn.Module):
    """x + 1/b * sin^2(a x), per channel (reference activations.py:62-120)."""

    def __init__(self, in_features, alpha=1.0, alpha_trainable=True, alpha_logscale=False):
        super().__init__()
        self.in_features = in_features
        self.alpha_logscale = alpha_logscale
        if self.alpha_logscale:
            self.alpha = Parameter(torch.zeros(in_features) * alpha)
            self.beta = Parameter(torch.zeros(in_features) * alpha)
        else:
            self.alpha = Parameter(torch.ones(in_features) * alpha)
            self.beta = Parameter(torch.ones(in_features) * alpha)
        self.alpha.requires_grad = alpha_trainable
        self.beta.requires_grad = alpha_trainable
        self.no_div_by_zero = 0.000000001

    def forward(self, x):
        raise NotImplementedError("SnakeBeta is a parameter holder here: wrap it in Activation1d (the fused sm_100a "
                                  "kernel); this package has no torch compute path")
