"""ECAPA-TDNN speaker encoder of the v1 generator (reference ``indextts/BigVGAN/ECAPA_TDNN.py:429-581`` with the layer
wrappers of ``indextts/BigVGAN/nnet/CNN.py:305-516`` and ``nnet/normalization.py``).

Host-side torch module, NOT part of the vocoder's kernel path: it runs once per call on the short reference mel
(``models.py:202``) and produces the ``[B, 1, D]`` embedding the generator is conditioned on.  Written against the
reference's state-dict layout (``blocks.N.conv.conv.weight``, ``...norm.norm.running_mean``, ``asp.tdnn...``, ``fc.conv...``)
so a v1 checkpoint loads with ``strict=True``; pinned by ``tests/golden/ecapa.npz`` (``make_golden_variants.py``).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class _SameConv1d(nn.Module):
    """``[N, C, L]`` conv with reflect 'same' padding (CNN.py:411-481 with ``skip_transpose=True``, stride 1:
    ``dilation * (k - 1) / 2`` samples each side)."""

    def __init__(self, in_channels, out_channels, kernel_size, dilation=1, groups=1):
        super().__init__()
        self.pad = dilation * (kernel_size - 1) // 2
        self.conv = nn.Conv1d(in_channels, out_channels, kernel_size, dilation=dilation, groups=groups)

    def forward(self, x):
        if self.pad:
            x = F.pad(x, (self.pad, self.pad), mode="reflect")
        return self.conv(x)


class _BatchNorm(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.norm = nn.BatchNorm1d(channels)

    def forward(self, x):
        return self.norm(x)


class TDNNBlock(nn.Module):
    """conv -> ReLU -> BatchNorm (ECAPA_TDNN.py:79-128)."""

    def __init__(self, in_channels, out_channels, kernel_size, dilation, groups=1):
        super().__init__()
        self.conv = _SameConv1d(in_channels, out_channels, kernel_size, dilation, groups)
        self.activation = nn.ReLU()
        self.norm = _BatchNorm(out_channels)

    def forward(self, x):
        return self.norm(self.activation(self.conv(x)))


class Res2NetBlock(nn.Module):
    """Hierarchical residual over ``scale`` channel chunks (ECAPA_TDNN.py:131-191)."""

    def __init__(self, in_channels, out_channels, scale=8, kernel_size=3, dilation=1):
        super().__init__()
        assert in_channels % scale == 0 and out_channels % scale == 0
        self.blocks = nn.ModuleList([TDNNBlock(in_channels // scale, out_channels // scale, kernel_size, dilation)
                                     for _ in range(scale - 1)])
        self.scale = scale

    def forward(self, x):
        outs, prev = [], None
        for i, chunk in enumerate(torch.chunk(x, self.scale, dim=1)):
            if i == 0:
                prev = chunk
            elif i == 1:
                prev = self.blocks[0](chunk)
            else:
                prev = self.blocks[i - 1](chunk + prev)
            outs.append(prev)
        return torch.cat(outs, dim=1)


def _length_mask(lengths, L, device):
    """``[N, 1, L]`` float mask of the first ``lengths * L`` positions (ECAPA_TDNN.py:16-62)."""
    steps = torch.arange(L, device=device, dtype=lengths.dtype).expand(len(lengths), L)
    return (steps < (lengths * L).unsqueeze(1)).to(lengths.dtype).unsqueeze(1)


class SEBlock(nn.Module):
    """Squeeze-and-excitation over time (ECAPA_TDNN.py:194-242)."""

    def __init__(self, in_channels, se_channels, out_channels):
        super().__init__()
        self.conv1 = _SameConv1d(in_channels, se_channels, 1)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = _SameConv1d(se_channels, out_channels, 1)
        self.sigmoid = nn.Sigmoid()

    def forward(self, x, lengths=None):
        if lengths is not None:
            mask = _length_mask(lengths, x.shape[-1], x.device)
            s = (x * mask).sum(dim=2, keepdim=True) / mask.sum(dim=2, keepdim=True)
        else:
            s = x.mean(dim=2, keepdim=True)
        return self.sigmoid(self.conv2(self.relu(self.conv1(s)))) * x


class AttentiveStatisticsPooling(nn.Module):
    """Attention-weighted mean and standard deviation over time (ECAPA_TDNN.py:245-338)."""

    def __init__(self, channels, attention_channels=128, global_context=True):
        super().__init__()
        self.eps = 1e-12
        self.global_context = global_context
        self.tdnn = TDNNBlock(channels * 3 if global_context else channels, attention_channels, 1, 1)
        self.tanh = nn.Tanh()
        self.conv = _SameConv1d(attention_channels, channels, 1)

    def _stats(self, x, w):
        mean = (w * x).sum(2)
        std = torch.sqrt((w * (x - mean.unsqueeze(2)).pow(2)).sum(2).clamp(self.eps))
        return mean, std

    def forward(self, x, lengths=None):
        L = x.shape[-1]
        if lengths is None:
            lengths = torch.ones(x.shape[0], device=x.device)
        mask = _length_mask(lengths, L, x.device)
        if self.global_context:
            mean, std = self._stats(x, mask / mask.sum(dim=2, keepdim=True).float())
            attn = torch.cat([x, mean.unsqueeze(2).repeat(1, 1, L), std.unsqueeze(2).repeat(1, 1, L)], dim=1)
        else:
            attn = x
        attn = self.conv(self.tanh(self.tdnn(attn)))
        attn = F.softmax(attn.masked_fill(mask == 0, float("-inf")), dim=2)
        mean, std = self._stats(x, attn)
        return torch.cat((mean, std), dim=1).unsqueeze(2)


class SERes2NetBlock(nn.Module):
    """TDNN -> Res2Net -> TDNN -> SE, plus the (projected) input (ECAPA_TDNN.py:341-426)."""

    def __init__(self, in_channels, out_channels, res2net_scale=8, se_channels=128, kernel_size=1, dilation=1, groups=1):
        super().__init__()
        self.out_channels = out_channels
        self.tdnn1 = TDNNBlock(in_channels, out_channels, 1, 1, groups)
        self.res2net_block = Res2NetBlock(out_channels, out_channels, res2net_scale, kernel_size, dilation)
        self.tdnn2 = TDNNBlock(out_channels, out_channels, 1, 1, groups)
        self.se_block = SEBlock(out_channels, se_channels, out_channels)
        self.shortcut = None
        if in_channels != out_channels:
            self.shortcut = _SameConv1d(in_channels, out_channels, 1)

    def forward(self, x, lengths=None):
        residual = self.shortcut(x) if self.shortcut is not None else x
        x = self.tdnn2(self.res2net_block(self.tdnn1(x)))
        return self.se_block(x, lengths) + residual


class ECAPA_TDNN(nn.Module):
    """``[B, T, input_size]`` features -> ``[B, 1, lin_neurons]`` embedding (ECAPA_TDNN.py:429-581)."""

    def __init__(self, input_size, device="cpu", lin_neurons=192, channels=(512, 512, 512, 512, 1536),
                 kernel_sizes=(5, 3, 3, 3, 1), dilations=(1, 2, 3, 4, 1), attention_channels=128, res2net_scale=8,
                 se_channels=128, global_context=True, groups=(1, 1, 1, 1, 1)):
        super().__init__()
        assert len(channels) == len(kernel_sizes) == len(dilations)
        self.channels = list(channels)
        self.blocks = nn.ModuleList([TDNNBlock(input_size, channels[0], kernel_sizes[0], dilations[0], groups[0])])
        for i in range(1, len(channels) - 1):
            self.blocks.append(SERes2NetBlock(channels[i - 1], channels[i], res2net_scale, se_channels, kernel_sizes[i],
                                              dilations[i], groups[i]))
        self.mfa = TDNNBlock(channels[-2] * (len(channels) - 2), channels[-1], kernel_sizes[-1], dilations[-1], groups[-1])
        self.asp = AttentiveStatisticsPooling(channels[-1], attention_channels, global_context)
        self.asp_bn = _BatchNorm(channels[-1] * 2)
        self.fc = _SameConv1d(channels[-1] * 2, lin_neurons, 1)

    def forward(self, x, lengths=None):
        x = x.transpose(1, 2)
        feats = []
        for i, layer in enumerate(self.blocks):
            x = layer(x) if i == 0 else layer(x, lengths=lengths)
            feats.append(x)
        x = self.mfa(torch.cat(feats[1:], dim=1))
        x = self.fc(self.asp_bn(self.asp(x, lengths=lengths)))
        return x.transpose(1, 2)
