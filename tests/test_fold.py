"""Time folding of small-channel convs (include/bvg.h, bvg_conv_desc::fold): the folded weight applied to the
[B, T/r, r*C] view must reproduce the reference Conv1d exactly (same products, same zero padding)."""
import pytest
import torch
import torch.nn.functional as F


@pytest.mark.parametrize("k,d", [(3, 1), (3, 3), (3, 5), (7, 1), (11, 1), (11, 3), (7, 5)])
@pytest.mark.parametrize("r", [2, 4])
def test_folded_weight_equals_reference_conv(k, d, r):
    from index_tts_b200.bigvgan import fold_conv1d_weight
    torch.manual_seed(k * 100 + d * 10 + r)
    C, T, B = 6, 8 * r, 2
    x = torch.randn(B, C, T, dtype=torch.float64)
    w = torch.randn(C, C, k, dtype=torch.float64)
    b = torch.randn(C, dtype=torch.float64)
    ref = F.conv1d(x, w, b, dilation=d, padding=(k - 1) // 2 * d)                  # reference bigvgan.py:56-88
    wf = fold_conv1d_weight(w, d, r)
    taps = wf.shape[2]
    assert taps == 2 * -(-((k - 1) // 2 * d) // r) + 1
    # [B, C, T] -> channels-last [B, T, C] -> view [B, T/r, r*C] -> channels-first for F.conv1d
    xf = x.permute(0, 2, 1).reshape(B, T // r, r * C).permute(0, 2, 1)
    yf = F.conv1d(xf, wf, b.repeat(r), padding=(taps - 1) // 2)
    y = yf.permute(0, 2, 1).reshape(B, T, C).permute(0, 2, 1)
    assert torch.allclose(y, ref, rtol=0, atol=1e-12)
