"""Whole-generator parity on a B200 through the reference-shaped module (BigVGAN.forward -> bvg_forward)."""
import os

import numpy as np
import pytest
import torch

from helpers import MINI_H, O, build_model, snr_db

pytestmark = pytest.mark.gpu


def _check_fp32(ref, wav, what):
    err = float(np.abs(wav - ref).max())
    snr = snr_db(ref, wav)
    print(f"{what}: fp32 max-abs {err:.3e}, SNR {snr:.1f} dB")
    assert err <= 1e-4, what       # BASELINE.json: fp32 max-abs waveform error <= 1e-4
    assert snr >= 80.0, what       # ... with SNR >= 80 dB


def _check_bf16(ref, wav, what):
    snr = snr_db(ref, wav)
    print(f"{what}: bf16 SNR {snr:.1f} dB, max-abs {float(np.abs(wav - ref).max()):.3e}")
    assert snr >= 40.0, what       # BASELINE.json: bf16 SNR >= 40 dB against the fp32 reference


@pytest.mark.parametrize("form", ["wn", "plain"])
def test_mini_generator_golden(golden_dir, form):
    """Mini generator (initial channel 512), B = 2, ragged T = 23, reference outputs from tests/golden."""
    g = np.load(os.path.join(golden_dir, f"mini_forward_{form}.npz"))
    sd = O.make_state_dict(MINI_H, seed=int(g["seed"]), weight_norm=(form == "wn"))
    m = build_model(MINI_H, sd, precision="fp32")
    mel = torch.from_numpy(g["mel"]).cuda()
    with torch.no_grad():
        wav = m(mel).cpu().numpy()
    assert wav.shape == g["wav"].shape
    _check_fp32(g["wav64"], wav, f"mini/{form}")
    m.set_precision("bf16")
    with torch.no_grad():
        wav = m(mel).cpu().numpy()
    _check_bf16(g["wav64"], wav, f"mini/{form}")
    if form == "wn":  # remove_weight_norm() afterwards must not change the result (re-pack path)
        m.remove_weight_norm()
        with torch.no_grad():
            wav2 = m(mel).cpu().numpy()
        # torch's fold and bvg_fold_weight_norm differ in the last ulp, which moves some bf16 weight roundings: the two
        # bf16 runs are equally far from the reference but not identical -- judge each against the reference.
        _check_bf16(g["wav64"], wav2, f"mini/{form} after remove_weight_norm")
        m.set_precision("fp32")
        with torch.no_grad():
            wav3 = m(mel).cpu().numpy()
        _check_fp32(g["wav64"], wav3, f"mini/{form} after remove_weight_norm")


@pytest.fixture(scope="module")
def full_models():
    sd = O.make_state_dict(O.DEFAULT_HPARAMS, seed=11)
    return sd, build_model(O.DEFAULT_HPARAMS, sd, precision="fp32")


def test_full_generator_golden(golden_dir, full_models):
    """bigvgan_v2_22khz_80band_256x (112 M parameters), reference outputs from tests/golden/full_forward.npz."""
    g = np.load(os.path.join(golden_dir, "full_forward.npz"))
    _, m = full_models
    for tag in ("a", "b"):
        mel = torch.from_numpy(g[f"mel_{tag}"]).cuda()
        m.set_precision("fp32")
        with torch.no_grad():
            wav = m(mel).cpu().numpy()
        _check_fp32(g[f"wav_{tag}"], wav, f"full/{tag}")
        m.set_precision("bf16")
        with torch.no_grad():
            wav = m(mel).cpu().numpy()
        _check_bf16(g[f"wav_{tag}"], wav, f"full/{tag}")


def test_forward_pcm16_equals_callers_conversion(golden_dir, full_models):
    """bvg_forward_pcm16 == infer_v2.py:866 + :912 applied to bvg_forward's waveform, both precisions, incl. the clamp."""
    g = np.load(os.path.join(golden_dir, "full_forward.npz"))
    _, m = full_models
    for prec in ("fp32", "bf16"):
        m.set_precision(prec)
        for scale in (1.0, 40.0):  # 40x: the random-init waveform (max 0.05) reaches the clamp
            mel = torch.from_numpy(g["mel_a"]).cuda()
            with torch.no_grad():
                w0 = m.conv_post.weight.detach().clone()
                m.conv_post.weight.mul_(scale)
                try:
                    wav = m(mel)
                    pcm = m.forward_pcm16(mel)
                finally:
                    m.conv_post.weight.copy_(w0)
            assert pcm.dtype == torch.int16 and pcm.shape == wav.shape
            assert torch.equal(pcm, torch.clamp(32767 * wav, -32767.0, 32767.0).type(torch.int16))
            if scale > 1.0:
                assert int(pcm.abs().max()) == 32767
    m.set_precision("fp32")
    assert m.forward_pcm16(torch.zeros(0, 80, 5, device="cuda")).shape == (0, 1, 1280)


def test_full_generator_clamp_path(golden_dir):
    """1.3x weights + log-mel-like input: 16 % of the samples hit the final clamp(-1, 1)."""
    g = np.load(os.path.join(golden_dir, "full_forward.npz"))
    sd = O.make_state_dict(O.DEFAULT_HPARAMS, seed=int(g["seed_c"]), scale=float(g["scale_c"]))
    m = build_model(O.DEFAULT_HPARAMS, sd, precision="fp32")
    with torch.no_grad():
        wav = m(torch.from_numpy(g["mel_c"]).cuda()).cpu().numpy()
    assert np.abs(wav).max() <= 1.0 and (np.abs(wav) == 1.0).mean() > 0.05
    # amplified weights amplify rounding too: the clamp path is checked at SNR >= 60 dB / max-abs 5e-3
    assert snr_db(g["wav_c"], wav) >= 60.0 and np.abs(wav - g["wav_c"]).max() < 5e-3
    m.set_precision("bf16")
    with torch.no_grad():
        wav = m(torch.from_numpy(g["mel_c"]).cuda()).cpu().numpy()
    assert np.abs(wav).max() <= 1.0 and snr_db(g["wav_c"], wav) >= 25.0


def test_cfg1_vs_oracle_and_module_contract(full_models):
    """BASELINE.json configs[0]: one 5 s mel (80 x 431), batch 1, against the CPU oracle run here on the host cores.
    Also: non-contiguous time-slice input, inference_mode input, output shape/dtype/range."""
    sd, m = full_models
    gen = torch.Generator().manual_seed(0)
    mel_full = torch.randn(1, 80, 431 + 20, generator=gen)
    mel = mel_full[:, :, 20:]                       # non-contiguous like infer_v2.py:856
    ref = O.forward(sd, mel.numpy(), None, np.float32)
    m.set_precision("fp32")
    with torch.inference_mode():
        wav = m(mel_full.cuda()[:, :, 20:])
    assert tuple(wav.shape) == (1, 1, 431 * 256) and wav.dtype == torch.float32
    assert float(wav.abs().max()) <= 1.0
    _check_fp32(ref, wav.cpu().numpy(), "cfg1")
    m.set_precision("bf16")
    with torch.no_grad():
        wav = m(mel.cuda()).cpu().numpy()
    _check_bf16(ref, wav, "cfg1")


def test_batch_items_are_independent_and_time_sharding(full_models):
    """B > 1 with independent edges; time-sharding with the 38-frame halo equals the unsharded forward."""
    _, m = full_models
    m.set_precision("fp32")
    gen = torch.Generator().manual_seed(5)
    mel = torch.randn(3, 80, 100, generator=gen).cuda()
    with torch.no_grad():
        wav = m(mel)
        one = m(mel[1:2])
    assert torch.allclose(wav[1:2], one, atol=1e-6)
    halo, hop = O.RECEPTIVE_HALO_FRAMES, 256
    with torch.no_grad():
        left = m(mel[:1, :, :50 + halo])[..., :50 * hop]
        right = m(mel[:1, :, 50 - halo:])[..., halo * hop:]
    stitched = torch.cat([left, right], dim=-1)
    assert float((stitched - wav[:1]).abs().max()) <= 1e-6


def test_empty_and_tiny_inputs(full_models):
    _, m = full_models
    m.set_precision("bf16")
    with torch.no_grad():
        assert tuple(m(torch.zeros(0, 80, 10, device="cuda")).shape) == (0, 1, 2560)
        assert tuple(m(torch.zeros(2, 80, 0, device="cuda")).shape) == (2, 1, 0)
        w = m(torch.randn(1, 80, 1, device="cuda"))
    assert tuple(w.shape) == (1, 1, 256) and torch.isfinite(w).all()
    with pytest.raises(ValueError):
        m(torch.zeros(1, 81, 4, device="cuda"))


def test_submodule_forwards_match_fused_plan(golden_dir):
    """Calling the reference-shaped submodules one by one ([B,C,T] tensors, like the reference's own forward loop)
    gives the same waveform as the fused bvg_forward plan."""
    g = np.load(os.path.join(golden_dir, "mini_forward_plain.npz"))
    sd = O.make_state_dict(MINI_H, seed=int(g["seed"]))
    m = build_model(MINI_H, sd, precision="fp32")
    mel = torch.from_numpy(g["mel"]).cuda()
    with torch.no_grad():
        x = m.conv_pre(mel)
        np.testing.assert_allclose(x.cpu().numpy(), g["conv_pre"], atol=2e-5)
        for i in range(m.num_upsamples):
            x = m.ups[i][0](x)
            if i == 0:
                np.testing.assert_allclose(x.cpu().numpy(), g["ups0"], atol=2e-5)
            xs = None
            for j in range(m.num_kernels):
                r = m.resblocks[i * m.num_kernels + j](x)
                if i == 0 and j == 0:
                    np.testing.assert_allclose(r.cpu().numpy(), g["rb0"], atol=5e-5)
                xs = r if xs is None else xs + r
            x = xs / m.num_kernels
        x = m.activation_post(x)
        full = m(mel)
    from index_tts_b200 import ops
    post = ops.conv_post(ops.ncl_to_nlc(x, torch.float32), m.conv_post.weight, None, False)
    assert float((post - full).abs().max()) <= 2e-6


def test_forward_on_side_stream_two_threads_and_graph_capture(full_models):
    """SURVEY.md 8(b) threading / stream contract: launches go to the caller's current stream, two host threads may call
    forward concurrently (serve_tars.py runs infer from a 2-worker ThreadPoolExecutor), and the plan (three-stream fork /
    join, PDL launches) is CUDA-graph capturable.  All results must equal the plain call bit for bit."""
    import threading
    sd, m = full_models
    m.set_precision("bf16")
    mels = [torch.randn(1, 80, 40 + 7 * i, generator=torch.Generator().manual_seed(30 + i)).cuda() for i in range(4)]
    with torch.no_grad():
        want = [m(x).clone() for x in mels]
    torch.cuda.synchronize()
    # side stream
    s = torch.cuda.Stream()
    with torch.cuda.stream(s), torch.no_grad():
        got = m(mels[0])
    s.synchronize()
    assert torch.equal(got, want[0])
    # two threads, each on its own stream, interleaved calls
    out, errs = {}, []

    def worker(tid):
        try:
            st = torch.cuda.Stream()
            with torch.cuda.stream(st), torch.no_grad():
                for rep in range(3):
                    for i in range(tid, 4, 2):
                        out[(tid, rep, i)] = m(mels[i])
            st.synchronize()
        except Exception as e:  # pragma: no cover
            errs.append(e)

    ts = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    for (tid, rep, i), y in out.items():
        assert torch.equal(y, want[i]), (tid, rep, i)
    # CUDA graph capture + replay
    static_in = mels[1].clone()
    g = torch.cuda.CUDAGraph()
    with torch.no_grad():
        m(static_in)  # warm (workspace allocation, plan)
        torch.cuda.synchronize()
        with torch.cuda.graph(g):
            static_out = m(static_in)
    for i in (3, 1):
        if mels[i].shape == static_in.shape:
            static_in.copy_(mels[i])
            g.replay()
            torch.cuda.synchronize()
            assert torch.equal(static_out, want[i])
    m.set_precision("fp32")


def test_torch_compile_dynamic(full_models):
    """infer_v2.py:267-270 wraps the vocoder in torch.compile(dynamic=True): the custom op + fake kernel must trace."""
    sd, m = full_models
    m.set_precision("bf16")
    cm = torch.compile(m, dynamic=True)
    with torch.no_grad():
        for T in (33, 47):
            x = torch.randn(1, 80, T, generator=torch.Generator().manual_seed(T)).cuda()
            assert torch.equal(cm(x), m(x))
    m.set_precision("fp32")


def test_cuda_graph_replay_cache_is_bit_identical(full_models):
    """``set_cuda_graphs``: small calls replayed from per-shape, per-thread CUDA graphs equal the eager plan bit for bit --
    repeated shapes, more shapes than cache entries, a strided mel, a side stream, two host threads, PCM output."""
    import threading
    sd, m = full_models
    m.set_precision("bf16")
    gens = [torch.Generator().manual_seed(70 + i) for i in range(7)]
    mels = [torch.randn(1 + (i % 2), 80, 24 + 5 * i, generator=gens[i]).cuda() for i in range(7)]
    with torch.no_grad():
        want = [m(x).clone() for x in mels]
        want_pcm = m.forward_pcm16(mels[0]).clone()
        m.set_cuda_graphs(4096)
        try:
            for rep in range(3):
                for i, x in enumerate(mels):  # 7 shapes > 4 cache entries: eviction and re-capture
                    assert torch.equal(m(x), want[i]), (rep, i)
            assert torch.equal(m.forward_pcm16(mels[0]), want_pcm)
            wide = torch.randn(1, 80, 60, generator=gens[0]).cuda()
            view = wide[:, :, 11:11 + mels[0].shape[2]]
            assert not view.is_contiguous()
            assert torch.equal(m(view), m.set_cuda_graphs(0)(view.contiguous()))
            m.set_cuda_graphs(4096)
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                a = m(mels[2])
                b = m(mels[2])
            c = m(mels[2])  # same graph from another stream of the same thread
            torch.cuda.synchronize()
            assert torch.equal(a, want[2]) and torch.equal(b, want[2]) and torch.equal(c, want[2])
            big = torch.randn(4, 80, 2000, generator=gens[1]).cuda()  # above the threshold: eager path
            assert m(big).shape == (4, 1, 2000 * 256)
            errs, out = [], {}

            def worker(tid):
                try:
                    st = torch.cuda.Stream()
                    with torch.cuda.stream(st), torch.no_grad():
                        for rep in range(4):
                            for i in range(tid, 4, 2):
                                out[(tid, rep, i)] = m(mels[i])
                    st.synchronize()
                except Exception:  # pragma: no cover
                    import traceback
                    errs.append(traceback.format_exc())

            ts = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
            [t.start() for t in ts]
            [t.join() for t in ts]
            assert not errs, "\n".join(errs)
            for (tid, rep, i), y in out.items():
                assert torch.equal(y, want[i]), (tid, rep, i)
        finally:
            m.set_cuda_graphs(0)
            m.set_precision("fp32")


def test_cfg2_full_batch_vs_oracle(full_models):
    """BASELINE.json configs[1] at full size (16 x 862 frames, the bench workload) through BigVGAN.forward: the
    oracle checks one utterance end to end; the other fifteen are checked through batch independence (every utterance of
    the batched call equals its own B = 1 call to 1e-6 in fp32 mode, and stays on the bf16 bar in bf16 mode)."""
    sd, m = full_models
    g = torch.Generator().manual_seed(21)
    mel = torch.randn(16, 80, 862, generator=g)
    mel_d = mel.cuda()
    m.set_precision("fp32")
    with torch.no_grad():
        wav32 = m(mel_d)
    assert tuple(wav32.shape) == (16, 1, 862 * 256)
    picks = [15]
    refs = {i: O.forward(sd, mel[i:i + 1].numpy(), None, np.float32) for i in picks}
    for i in picks:
        _check_fp32(refs[i], wav32[i:i + 1].cpu().numpy(), f"cfg2 utterance {i}")
    with torch.no_grad():
        for i in range(16):
            if i in picks:
                continue
            one = m(mel_d[i:i + 1])
            assert float((one[0] - wav32[i]).abs().max()) <= 1e-6, f"utterance {i}: batched fp32 result differs from B = 1"
    m.set_precision("bf16")
    with torch.no_grad():
        wav16 = m(mel_d)
    for i in picks:
        _check_bf16(refs[i], wav16[i:i + 1].float().cpu().numpy(), f"cfg2 utterance {i}")
    ref32 = wav32.cpu().numpy().astype(np.float64)
    got16 = wav16.float().cpu().numpy()
    for i in range(16):
        assert snr_db(ref32[i], got16[i]) >= 40.0, f"utterance {i} in bf16 mode"
    m.set_precision("fp32")


def test_dynamic_tile_scheduler_opt_in_is_bit_identical(tmp_path):
    """BVG_CONV_DYNAMIC=1 (conv CTAs take tiles from a per-launch counter in the plan's workspace instead of static
    striding; off by default, read once per process): the mini generator's waveform in both precisions, ragged T and
    B > 1, equals the default scheduler's bit for bit (a tile computed twice or skipped would show)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = f"""
import sys
sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'tests')!r})
import numpy as np, torch
from helpers import MINI_H, O, build_model
sd = O.make_state_dict(MINI_H, seed=9)
m = build_model(MINI_H, sd, precision="bf16")
g = torch.Generator().manual_seed(2)
out = {{}}
for name, (B, T) in (("a", (3, 301)), ("b", (1, 64))):
    mel = torch.randn(B, 80, T, generator=g).cuda()
    for prec in ("bf16", "fp32"):
        m.set_precision(prec)
        with torch.no_grad():
            out[name + prec] = m(mel).cpu().numpy()
np.savez(sys.argv[1], **out)
"""
    files = []
    for dyn in ("0", "1"):
        f = str(tmp_path / f"dyn{dyn}.npz")
        env = dict(os.environ, BVG_CONV_DYNAMIC=dyn)
        r = subprocess.run([sys.executable, "-c", code, f], env=env, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-3000:]
        files.append(np.load(f))
    for k in files[0].files:
        assert np.isfinite(files[1][k]).all()
        assert np.array_equal(files[0][k], files[1][k]), f"{k}: dynamic scheduler changed the result"
