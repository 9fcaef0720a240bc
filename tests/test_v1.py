"""The v1 conditioned generator (reference indextts/BigVGAN/models.py), the AMPBlock2 / Snake configurations of the fused
launch plan, and the ECAPA-TDNN speaker encoder, against fixtures produced by the reference itself
(tests/golden/make_golden_variants.py).  CPU tests cover the host logic; ``-m gpu`` tests call through the C ABI."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import bigvgan_oracle as O
from helpers import V1_MINI_H, build_model, build_seeded_ecapa

VARIANTS = ["v1", "v1_amp2", "v1_amp2_snake", "v1_amp1_snake_nocond"]


class _H(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def _variants(golden_dir):
    return np.load(os.path.join(golden_dir, "variants.npz"))


def test_ecapa_matches_reference(golden_dir):
    from index_tts_b200 import ecapa
    g = np.load(os.path.join(golden_dir, "ecapa.npz"))
    m = build_seeded_ecapa(ecapa)
    with torch.no_grad():
        e = m(torch.from_numpy(g["mel_ref"])).numpy()
        el = m(torch.from_numpy(g["mel_ref"]), torch.from_numpy(g["lens"])).numpy()
    assert e.shape == g["emb"].shape == (3, 1, 48)
    np.testing.assert_allclose(e, g["emb"], rtol=0, atol=2e-5)
    np.testing.assert_allclose(el, g["emb_lens"], rtol=0, atol=2e-5)
    assert np.abs(g["emb"] - g["emb_lens"]).max() > 1e-3  # the fixture exercises the length masks


def test_v1_state_dict_layout_matches_reference(golden_dir):
    from index_tts_b200 import v1
    layout = json.loads(str(_variants(golden_dir)["v1_state_dict_layout"]))
    m = v1.BigVGAN(_H(dict(V1_MINI_H)), use_cuda_kernel=False)
    ours = {k: list(v.shape) for k, v in m.state_dict().items()}
    assert ours == layout
    assert m.h["use_cuda_kernel"] is False
    m.remove_weight_norm()
    assert "conv_pre.weight" in m.state_dict() and "conv_pre.weight_g" not in m.state_dict()


def test_v1_errors_match_reference():
    from index_tts_b200 import v1, _lib
    with pytest.raises(NotImplementedError):
        v1.BigVGAN(_H(dict(V1_MINI_H, activation="relu")))
    m = v1.BigVGAN(_H(dict(V1_MINI_H, upsample_initial_channel=64)))
    with pytest.raises(_lib.BvgError):  # no CPU path
        m.forward_with_embedding(torch.zeros(1, 4, 64), torch.zeros(1, 32))
    # v1 AMPBlock2 keeps two convs whatever the tuple lists (models.py:79-93)
    assert len(v1.AMPBlock2(_H(dict(V1_MINI_H)), 16, 3, (1, 3, 5), activation="snake").convs) == 2


def _build_v1(h, sd_np, precision):
    from index_tts_b200 import v1
    m = v1.BigVGAN(_H(dict(h)), use_cuda_kernel=True)
    m.remove_weight_norm()
    missing, unexpected = m.load_state_dict({k: torch.from_numpy(v) for k, v in sd_np.items()}, strict=False)
    assert not unexpected and all(k.startswith("speaker_encoder.") for k in missing)
    return m.to("cuda").eval().set_precision(precision)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", VARIANTS)
def test_gpu_v1_generator_variants_fp32(golden_dir, tag):
    g = _variants(golden_dir)
    h = json.loads(str(g[f"{tag}_h"]))
    m = _build_v1(h, O.make_state_dict(h, seed=int(g[f"{tag}_seed"])), "fp32")
    with torch.no_grad():
        wav = m.forward_with_embedding(torch.from_numpy(g[f"{tag}_x"]).cuda(), torch.from_numpy(g[f"{tag}_emb"]).cuda())
    ref = g[f"{tag}_wav"]
    wav = wav.cpu().numpy()
    assert wav.shape == ref.shape
    # BASELINE.json tolerance for the fp32 mode: max-abs <= 1e-4 and SNR >= 80 dB
    assert np.abs(wav - ref).max() <= 1e-4 and O.snr_db(ref, wav) >= 80.0, (np.abs(wav - ref).max(), O.snr_db(ref, wav))


@pytest.mark.gpu
@pytest.mark.parametrize("tag", VARIANTS)
def test_gpu_v1_generator_variants_bf16(golden_dir, tag):
    g = _variants(golden_dir)
    h = json.loads(str(g[f"{tag}_h"]))
    m = _build_v1(h, O.make_state_dict(h, seed=int(g[f"{tag}_seed"])), "bf16")
    with torch.no_grad():
        wav = m.forward_with_embedding(torch.from_numpy(g[f"{tag}_x"]).cuda(), torch.from_numpy(g[f"{tag}_emb"]).cuda())
    ref = g[f"{tag}_wav"]
    # BASELINE.json tolerance for the bf16 mode: SNR >= 40 dB against the fp32 reference
    assert O.snr_db(ref, wav.cpu().numpy()) >= 40.0, O.snr_db(ref, wav.cpu().numpy())


@pytest.mark.gpu
def test_gpu_v1_forward_with_speaker_encoder_and_batch_independence():
    """``forward(x, mel_ref)`` = speaker encoder + decoder; every utterance gets ITS OWN conditioning vector."""
    h = dict(V1_MINI_H)
    m = _build_v1(h, O.make_state_dict(h, seed=31), "fp32")
    gen = torch.Generator().manual_seed(5)
    x = torch.randn(3, 6, h["gpt_dim"], generator=gen).cuda()
    mel_ref = torch.randn(3, 40, h["num_mels"], generator=gen).cuda()
    with torch.no_grad():
        wav, loss = m(x, mel_ref)
        emb = m.speaker_encoder(mel_ref)
        dec = m.forward_with_embedding(x, emb)
        one = m.forward_with_embedding(x[1:2], emb[1:2])
    assert loss is None and wav.shape == (3, 1, 6 * 1024)
    sd = {k: v.detach().cpu().numpy() for k, v in m.state_dict().items() if not k.startswith("speaker_encoder.")}
    ref = O.forward_v1(sd, x.cpu().numpy(), emb[:, 0, :].cpu().numpy(), h, np.float64)
    assert np.abs(wav.cpu().numpy() - ref).max() <= 1e-4 and O.snr_db(ref, wav.cpu().numpy()) >= 80.0
    # the decoder treats utterances independently, bit for bit (the torch/cuDNN speaker encoder picks batch-dependent
    # algorithms, so the comparison is made with the embeddings given)
    assert torch.equal(one[0], dec[1])
    with pytest.raises(AttributeError):  # the contrastive branch reads the commented-out logit_scale (models.py:199, 205-207)
        m(x[:1], mel_ref[:2])


@pytest.mark.gpu
def test_gpu_cond_kernels_match_oracle():
    from index_tts_b200 import ops
    rng = np.random.default_rng(3)
    e = rng.standard_normal((5, 512)).astype(np.float32)
    w = rng.standard_normal((200, 512, 1)).astype(np.float32) / 16
    b = rng.standard_normal(200).astype(np.float32)
    y = ops.cond_linear(torch.from_numpy(e).cuda(), torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda()).cpu().numpy()
    ref = e.astype(np.float64) @ w[:, :, 0].astype(np.float64).T + b
    np.testing.assert_allclose(y, ref, rtol=0, atol=2e-5)
    for T, C in ((1, 8), (37, 24), (1000, 1536)):
        x = rng.standard_normal((3, T, C)).astype(np.float32)
        v = rng.standard_normal((3, C)).astype(np.float32)
        xt, vt = torch.from_numpy(x).cuda(), torch.from_numpy(v).cuda()
        got = ops.add_channel_vector_nlc(xt, vt, torch.float32).cpu().numpy()
        assert np.array_equal(got, x + v[:, None, :])  # one fp32 add per element: bit-exact
        got16 = ops.add_channel_vector_nlc(xt, vt, torch.bfloat16)
        assert torch.equal(got16.cpu(), torch.from_numpy(x + v[:, None, :]).bfloat16())
        assert torch.equal(ops.add_channel_vector_nlc(xt.bfloat16(), None, torch.float32).cpu(), xt.bfloat16().float().cpu())


@pytest.mark.gpu
@pytest.mark.parametrize("activation", ["snakebeta", "snake"])
def test_gpu_v2_generator_with_ampblock2_plan(activation):
    """``resblock "2"`` through the fused v2 launch plan (bvg_model_desc::resblock = 2) against the oracle."""
    h = dict(O.DEFAULT_HPARAMS, resblock="2", activation=activation, upsample_initial_channel=512)
    sd = O.make_state_dict(h, seed=41)
    mel = np.random.default_rng(8).standard_normal((2, 80, 21)).astype(np.float32)
    ref = O.forward(sd, mel, h, np.float64)
    m = build_model(h, sd, precision="fp32")
    with torch.no_grad():
        wav = m(torch.from_numpy(mel).cuda()).cpu().numpy()
    assert np.abs(wav - ref).max() <= 1e-4 and O.snr_db(ref, wav) >= 80.0
    m.set_precision("bf16")
    with torch.no_grad():
        wav16 = m(torch.from_numpy(mel).cuda()).cpu().numpy()
    assert O.snr_db(ref, wav16) >= 40.0


@pytest.mark.gpu
def test_gpu_v1_unfused_conditioning_path_matches_too(golden_dir):
    """BVG_COND_FUSED=0 (separate add kernels instead of the conv-epilogue bias; read once per process, hence the
    subprocess) gives the same parity against the reference fixture."""
    import subprocess
    import sys
    code = (
        "import sys, json, numpy as np, torch\n"
        f"sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})\n"
        f"sys.path.insert(0, {os.path.dirname(os.path.abspath(__file__))!r})\n"
        "import test_v1 as t\n"
        "from oracle import bigvgan_oracle as O\n"
        f"g = np.load({os.path.join(golden_dir, 'variants.npz')!r})\n"
        "h = json.loads(str(g['v1_h']))\n"
        "for prec, bar in (('fp32', 80.0), ('bf16', 40.0)):\n"
        "    m = t._build_v1(h, O.make_state_dict(h, seed=int(g['v1_seed'])), prec)\n"
        "    with torch.no_grad():\n"
        "        w = m.forward_with_embedding(torch.from_numpy(g['v1_x']).cuda(), torch.from_numpy(g['v1_emb']).cuda())\n"
        "    snr = O.snr_db(g['v1_wav'], w.cpu().numpy())\n"
        "    assert snr >= bar, (prec, snr)\n"
        "print('ok')\n")
    env = dict(os.environ, BVG_COND_FUSED="0")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-2000:]


@pytest.mark.gpu
def test_gpu_v1_edge_shapes():
    """Empty batch, empty time axis, a single latent frame (every conv sees only padding around it), fp32 and bf16."""
    h = dict(V1_MINI_H)
    sd = O.make_state_dict(h, seed=33)
    m = _build_v1(h, sd, "fp32")
    D, E = h["gpt_dim"], h["speaker_embedding_dim"]
    with torch.no_grad():
        assert m.forward_with_embedding(torch.zeros(0, 5, D).cuda(), torch.zeros(0, E).cuda()).shape == (0, 1, 5 * 1024)
        assert m.forward_with_embedding(torch.zeros(2, 0, D).cuda(), torch.zeros(2, E).cuda()).shape == (2, 1, 0)
        rng = np.random.default_rng(2)
        x = rng.standard_normal((2, 1, D)).astype(np.float32)
        e = rng.standard_normal((2, E)).astype(np.float32)
        ref = O.forward_v1(sd, x, e, h, np.float64)
        w = m.forward_with_embedding(torch.from_numpy(x).cuda(), torch.from_numpy(e).cuda()).cpu().numpy()
        assert w.shape == ref.shape == (2, 1, 1024)
        assert np.abs(w - ref).max() <= 1e-4 and O.snr_db(ref, w) >= 80.0
        m.set_precision("bf16")
        w16 = m.forward_with_embedding(torch.from_numpy(x).cuda(), torch.from_numpy(e).cuda()).cpu().numpy()
        assert O.snr_db(ref, w16) >= 40.0
        with pytest.raises(ValueError):
            m.forward_with_embedding(torch.zeros(2, 3, D + 8).cuda(), torch.zeros(2, E).cuda())
        with pytest.raises(ValueError):
            m.forward_with_embedding(torch.zeros(2, 3, D).cuda(), torch.zeros(3, E).cuda())
