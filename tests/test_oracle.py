"""The CPU oracle (oracle/) against the committed golden vectors produced by the reference itself
(tests/golden/make_golden.py).  Runs anywhere, no GPU, no /root/reference."""
import os

import numpy as np
import pytest

from oracle import bigvgan_oracle as O

MINI_H = dict(O.DEFAULT_HPARAMS, upsample_initial_channel=512)


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_filter_taps_match_reference(golden_dir):
    g = _load(golden_dir, "filter.npz")
    f = O.kaiser_sinc_filter1d(0.25, 0.3, 12)
    np.testing.assert_allclose(f, g["up"], rtol=0, atol=6e-8)
    np.testing.assert_allclose(f, g["down"], rtol=0, atol=6e-8)
    assert abs(f.sum() - 1.0) < 1e-12 and np.allclose(f, f[::-1])
    # SURVEY.md §8a row a9 values
    np.testing.assert_allclose(f[:6], [0.0020289647, 0.0093894657, -0.0255434588, -0.0576573834, 0.1285725832,
                                       0.4432097971], rtol=0, atol=6e-8)


def test_activation1d_matches_reference_all_lengths(golden_dir):
    g = _load(golden_dir, "act1d.npz")
    filt = _load(golden_dir, "filter.npz")["up"]
    for i in range(int(g["n"])):
        x, al, be = g[f"x{i}"], g[f"alpha{i}"], g[f"beta{i}"]
        y64 = O.activation1d(x.astype(np.float64), al, be, True, filt.astype(np.float64))
        np.testing.assert_allclose(y64, g[f"y64_{i}"], rtol=0, atol=1e-12)
        y32 = O.activation1d(x, al, be, True, filt)
        np.testing.assert_allclose(y32, g[f"y32_{i}"], rtol=0, atol=2e-5)
        assert y32.dtype == np.float32 and y32.shape == x.shape


def test_conv_layers_match_reference(golden_dir):
    g = _load(golden_dir, "convs.npz")
    for i in range(int(g["n_conv"])):
        y = O.conv1d(g[f"c_x{i}"].astype(np.float64), g[f"c_w{i}"].astype(np.float64), g[f"c_b{i}"],
                     dilation=int(g[f"c_d{i}"]))
        np.testing.assert_allclose(y, g[f"c_y{i}"], rtol=0, atol=1e-12)
    for i in range(int(g["n_convt"])):
        x, w, b, u = g[f"t_x{i}"].astype(np.float64), g[f"t_w{i}"].astype(np.float64), g[f"t_b{i}"], int(g[f"t_u{i}"])
        k = w.shape[-1]
        y = O.conv_transpose1d(x, w, b, stride=u, padding=(k - u) // 2)
        np.testing.assert_allclose(y, g[f"t_y{i}"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(O.conv_transpose1d_scatter(x, w, b, u, (k - u) // 2), y, atol=1e-12)
        assert y.shape[-1] == u * x.shape[-1]


def test_fold_weight_norm_definition():
    rng = np.random.default_rng(0)
    v = rng.standard_normal((6, 4, 3)).astype(np.float32)
    g = rng.uniform(0.5, 2, (6, 1, 1)).astype(np.float32)
    w = O.fold_weight_norm(v, g)
    np.testing.assert_allclose(np.sqrt((w.astype(np.float64) ** 2).sum(axis=(1, 2))), g.reshape(-1), rtol=1e-6)


@pytest.mark.parametrize("form", ["wn", "plain"])
def test_mini_generator_matches_reference(golden_dir, form):
    g = _load(golden_dir, f"mini_forward_{form}.npz")
    sd = O.make_state_dict(MINI_H, seed=int(g["seed"]), weight_norm=(form == "wn"))
    taps = {}
    wav = O.forward(sd, g["mel"], MINI_H, np.float64, taps)
    assert wav.shape == g["wav"].shape == (2, 1, 23 * 256)
    np.testing.assert_allclose(wav, g["wav64"], rtol=0, atol=1e-10 if form == "plain" else 1e-7)
    assert O.snr_db(g["wav64"], wav) > (200 if form == "plain" else 120)
    np.testing.assert_allclose(taps["conv_pre"], g["conv_pre"], atol=2e-5)
    np.testing.assert_allclose(taps["ups.0"], g["ups0"], atol=2e-5)
    np.testing.assert_allclose(taps["ups.3"][:, :, :512], g["ups3"], atol=2e-5)
    np.testing.assert_allclose(taps["conv_post"], g["conv_post"], atol=2e-5)
    wav32 = O.forward(sd, g["mel"], MINI_H, np.float32)
    assert np.abs(wav32 - g["wav"]).max() < 1e-5 and O.snr_db(g["wav"], wav32) > 100


@pytest.mark.slow
def test_full_generator_matches_reference(golden_dir):
    g = _load(golden_dir, "full_forward.npz")
    sd = O.make_state_dict(O.DEFAULT_HPARAMS, seed=int(g["seed"]))
    for tag in ("a", "b"):
        wav = O.forward(sd, g[f"mel_{tag}"], None, np.float32)
        assert wav.shape == g[f"wav_{tag}"].shape
        assert np.abs(wav - g[f"wav_{tag}"]).max() < 1e-5
        assert O.snr_db(g[f"wav_{tag}"], wav) > 100
    sdc = O.make_state_dict(O.DEFAULT_HPARAMS, seed=int(g["seed_c"]), scale=float(g["scale_c"]))
    wav = O.forward(sdc, g["mel_c"], None, np.float32)
    assert (np.abs(g["wav_c"]) >= 1.0).mean() > 0.05  # the fixture exercises the final clamp
    assert O.snr_db(g["wav_c"], wav) > 80


def test_time_sharding_with_halo_equals_unsharded():
    sd = O.make_state_dict(MINI_H, seed=3)
    mel = np.random.default_rng(5).standard_normal((1, 80, 100)).astype(np.float32)
    full = O.forward(sd, mel, MINI_H, np.float32)
    for n in (2, 3):
        sh = O.forward_time_sharded(sd, mel, n, MINI_H, np.float32)
        assert sh.shape == full.shape
        assert np.abs(sh - full).max() <= 1e-6
    # too small a halo must be visible (guards against a vacuous test)
    bad = O.forward_time_sharded(sd, mel, 2, MINI_H, np.float32, halo=2)
    assert np.abs(bad - full).max() > 1e-6


VARIANTS = ["v1", "v1_amp2", "v1_amp2_snake", "v1_amp1_snake_nocond"]


@pytest.mark.parametrize("tag", VARIANTS)
def test_v1_generator_variants_match_reference(golden_dir, tag):
    """The conditioned v1 generator (indextts/BigVGAN/models.py:201-258 after the speaker encoder) and the AMPBlock2 /
    Snake configurations, against outputs of the reference itself (tests/golden/make_golden_variants.py)."""
    import json
    g = _load(golden_dir, "variants.npz")
    h = json.loads(str(g[f"{tag}_h"]))
    sd = O.make_state_dict(h, seed=int(g[f"{tag}_seed"]))
    wav = O.forward_v1(sd, g[f"{tag}_x"], g[f"{tag}_emb"], h, np.float32)
    ref = g[f"{tag}_wav"]
    assert wav.shape == ref.shape
    assert np.abs(wav - ref).max() < 2e-5, np.abs(wav - ref).max()
    assert O.snr_db(ref, wav) > 100
