"""Generate the golden fixtures in this directory from the REFERENCE implementation.

Run in the authoring container only (needs ``/root/reference``):

    python tests/golden/make_golden.py

It imports the reference's own torch path unmodified (import-only stubs for matplotlib/librosa,
SURVEY.md Appendix A), feeds it deterministic inputs and stores inputs + outputs as small ``.npz``
files.  Weights that are too large to commit are regenerated from a NumPy seed by
``oracle.make_state_dict`` (PCG64, stable across machines) and loaded into the reference with
``load_state_dict(strict=True)``.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import bigvgan_oracle as O  # noqa: E402

MINI_H = dict(O.DEFAULT_HPARAMS, upsample_initial_channel=512)


def import_reference():
    for name in ["matplotlib", "matplotlib.pylab", "librosa", "librosa.util", "librosa.filters"]:
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.modules["librosa.util"].normalize = None
    sys.modules["librosa.filters"].mel = None
    sys.path.insert(0, "/root/reference")
    from indextts.s2mel.modules.bigvgan import bigvgan, activations
    from indextts.s2mel.modules.bigvgan.alias_free_activation.torch.act import Activation1d
    return bigvgan, activations, Activation1d


def ref_model(bigvgan, h, sd_np, weight_norm):
    hh = bigvgan.AttrDict(dict(h))
    m = bigvgan.BigVGAN(hh, use_cuda_kernel=False)
    if not weight_norm:
        m.remove_weight_norm()
    m.load_state_dict({k: torch.from_numpy(v) for k, v in sd_np.items()}, strict=True)
    if weight_norm:
        m.remove_weight_norm()
    return m.eval()


def main():
    torch.set_num_threads(os.cpu_count())
    bigvgan, activations, Activation1d = import_reference()
    rng = np.random.default_rng(20261104)

    # 1. filter taps (AF/torch/filter.py:30-62 via UpSample1d / DownSample1d buffers)
    a = Activation1d(activation=activations.SnakeBeta(8, alpha_logscale=True))
    np.savez(os.path.join(HERE, "filter.npz"),
             up=a.upsample.filter.numpy().reshape(-1), down=a.downsample.lowpass.filter.numpy().reshape(-1))

    # 2. Activation1d on ragged / tiny / odd lengths, random alpha/beta, fp32 and fp64
    cases = {}
    for idx, (B, C, T) in enumerate([(2, 8, 37), (1, 8, 1), (1, 16, 2), (3, 8, 5), (1, 8, 11), (1, 8, 12),
                                     (1, 24, 257), (2, 48, 300)]):
        x = rng.standard_normal((B, C, T)).astype(np.float32) * 2.0
        al = (0.5 * rng.standard_normal(C)).astype(np.float32)
        be = (0.5 * rng.standard_normal(C)).astype(np.float32)
        act = Activation1d(activation=activations.SnakeBeta(C, alpha_logscale=True))
        with torch.no_grad():
            act.act.alpha.copy_(torch.from_numpy(al))
            act.act.beta.copy_(torch.from_numpy(be))
            y32 = act(torch.from_numpy(x)).numpy()
            y64 = act.double()(torch.from_numpy(x).double()).numpy()
        cases.update({f"x{idx}": x, f"alpha{idx}": al, f"beta{idx}": be, f"y32_{idx}": y32, f"y64_{idx}": y64})
    cases["n"] = np.array(8)
    np.savez(os.path.join(HERE, "act1d.npz"), **cases)

    # 3. conv / transposed-conv layer cases with the reference's parameterisation
    cc = {}
    F = torch.nn.functional
    for idx, (B, Cin, Cout, T, k, d) in enumerate([(2, 8, 16, 33, 3, 1), (1, 16, 8, 40, 7, 3), (1, 8, 8, 64, 11, 5),
                                                   (1, 80, 32, 9, 7, 1), (2, 24, 1, 50, 7, 1), (1, 8, 8, 4, 11, 5)]):
        x = rng.standard_normal((B, Cin, T)).astype(np.float32)
        w = rng.standard_normal((Cout, Cin, k)).astype(np.float32) * 0.1
        b = rng.standard_normal(Cout).astype(np.float32)
        y = F.conv1d(torch.from_numpy(x).double(), torch.from_numpy(w).double(), torch.from_numpy(b).double(),
                     dilation=d, padding=O.get_padding(k, d)).numpy()
        cc.update({f"c_x{idx}": x, f"c_w{idx}": w, f"c_b{idx}": b, f"c_y{idx}": y, f"c_d{idx}": np.array(d)})
    for idx, (B, Cin, Cout, T, k, u) in enumerate([(2, 16, 8, 13, 8, 4), (1, 8, 8, 21, 4, 2), (1, 32, 16, 1, 8, 4),
                                                   (1, 16, 8, 2, 4, 2)]):
        x = rng.standard_normal((B, Cin, T)).astype(np.float32)
        w = rng.standard_normal((Cin, Cout, k)).astype(np.float32) * 0.1
        b = rng.standard_normal(Cout).astype(np.float32)
        y = F.conv_transpose1d(torch.from_numpy(x).double(), torch.from_numpy(w).double(),
                               torch.from_numpy(b).double(), stride=u, padding=(k - u) // 2).numpy()
        cc.update({f"t_x{idx}": x, f"t_w{idx}": w, f"t_b{idx}": b, f"t_y{idx}": y, f"t_u{idx}": np.array(u)})
    cc["n_conv"] = np.array(6)
    cc["n_convt"] = np.array(4)
    np.savez(os.path.join(HERE, "convs.npz"), **cc)

    # 4. mini generator (upsample_initial_channel 512), weight-norm form, B=2, ragged T
    for wn in (True, False):
        sd = O.make_state_dict(MINI_H, seed=7, weight_norm=wn)
        m = ref_model(bigvgan, MINI_H, sd, wn)
        mel = rng.standard_normal((2, 80, 23)).astype(np.float32)
        feats = {}
        hooks = []
        for name in ["conv_pre", "ups.0.0", "ups.3.0", "resblocks.0", "resblocks.10", "conv_post"]:
            mod = dict(m.named_modules())[name]
            hooks.append(mod.register_forward_hook(lambda _m, _i, o, name=name: feats.__setitem__(name, o.detach().numpy().copy())))
        with torch.no_grad():
            wav = m(torch.from_numpy(mel)).numpy()
            wav64 = m.double()(torch.from_numpy(mel).double()).numpy()
        for hk in hooks:
            hk.remove()
        np.savez(os.path.join(HERE, f"mini_forward_{'wn' if wn else 'plain'}.npz"), mel=mel, wav=wav, wav64=wav64,
                 seed=np.array(7), conv_pre=feats["conv_pre"], ups0=feats["ups.0.0"], ups3=feats["ups.3.0"][:, :, :512],
                 rb0=feats["resblocks.0"], rb10=feats["resblocks.10"][:, :, :512], conv_post=feats["conv_post"])

    # 5. full-size generator (bigvgan_v2_22khz_80band_256x), NumPy-seeded weights, short mels
    sd = O.make_state_dict(O.DEFAULT_HPARAMS, seed=11, weight_norm=False)
    m = ref_model(bigvgan, O.DEFAULT_HPARAMS, sd, False)
    out = {"seed": np.array(11)}
    for tag, (B, T) in {"a": (1, 40), "b": (2, 17)}.items():
        mel = rng.standard_normal((B, 80, T)).astype(np.float32)
        with torch.no_grad():
            wav = m(torch.from_numpy(mel)).numpy()
        out[f"mel_{tag}"] = mel
        out[f"wav_{tag}"] = wav
    # log-mel-like distribution (SURVEY.md §8d cfg1 secondary input) with 1.3x weights -> exercises the clamp
    sd3 = O.make_state_dict(O.DEFAULT_HPARAMS, seed=12, weight_norm=False, scale=1.3)
    m3 = ref_model(bigvgan, O.DEFAULT_HPARAMS, sd3, False)
    mel = np.clip(-5.0 + 2.0 * rng.standard_normal((1, 80, 24)), -11.5, 2.0).astype(np.float32)
    with torch.no_grad():
        wav = m3(torch.from_numpy(mel)).numpy()
    out["mel_c"] = mel
    out["wav_c"] = wav
    out["seed_c"] = np.array(12)
    out["scale_c"] = np.array(1.3)
    print("clamped fraction in wav_c:", float((np.abs(wav) >= 1.0).mean()), "rms", float(np.sqrt((wav ** 2).mean())))
    np.savez(os.path.join(HERE, "full_forward.npz"), **out)
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
