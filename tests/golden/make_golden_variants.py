"""Golden fixtures for the generator VARIANTS (AMPBlock2 / Snake configs, the v1 conditioned generator), generated from
the REFERENCE implementation.  Run in the authoring container only (needs ``/root/reference``):

    python tests/golden/make_golden_variants.py        ->  tests/golden/variants.npz

* v1 generator: ``indextts/BigVGAN/models.py`` ``BigVGAN`` run unmodified, ECAPA-TDNN speaker encoder included; the
  speaker embedding it computes (:202) is captured with a forward hook and stored, so the decoder half (:221-258) can be
  checked on its own.
* ``resblock "2"`` / ``activation "snake"`` run through the same class (its AMPBlock2 returns ``x``, models.py:117-123;
  the IndexTTS2 copy under ``s2mel/modules/bigvgan`` forgets to, bigvgan.py:232-236, and cannot run a "2" config at all).

Weights come from ``oracle.make_state_dict`` (NumPy PCG64 seeds) and are loaded with ``load_state_dict``.
"""
import json
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import bigvgan_oracle as O  # noqa: E402
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import V1_MINI_H, build_seeded_ecapa  # noqa: E402



class _H(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def main():
    for name in ["matplotlib", "matplotlib.pylab", "librosa", "librosa.util", "librosa.filters", "torchaudio"]:
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["matplotlib"].use = lambda *a, **k: None
    sys.modules["librosa.util"].normalize = None
    sys.modules["librosa.filters"].mel = None
    sys.path.insert(0, "/root/reference")
    from indextts.BigVGAN import models as ref_v1

    torch.set_num_threads(os.cpu_count())
    rng = np.random.default_rng(20261105)
    out = {}

    # v1 generator: latent [B, T, gpt_dim] + reference mel [B, T_ref, num_mels] -> wav.  The speaker embedding the
    # reference computes (models.py:202) is captured with a hook.  Variants: AMPBlock1/SnakeBeta (the published shape
    # class, scaled down), AMPBlock2/SnakeBeta, AMPBlock2/Snake, AMPBlock1/Snake without the per-stage conditioning.
    variants = {
        "v1": dict(V1_MINI_H),
        "v1_amp2": dict(V1_MINI_H, resblock="2", upsample_initial_channel=256, resblock_dilation_sizes=[[1, 3]] * 3,
                        upsample_rates=[4, 4, 2, 2], upsample_kernel_sizes=[8, 8, 4, 4]),
        "v1_amp2_snake": dict(V1_MINI_H, resblock="2", activation="snake", upsample_initial_channel=256,
                              resblock_dilation_sizes=[[1, 3]] * 3,  # models.py:79-93 hard-codes two convs
                              upsample_rates=[4, 2, 2], upsample_kernel_sizes=[8, 4, 4]),
        "v1_amp1_snake_nocond": dict(V1_MINI_H, activation="snake", upsample_initial_channel=128,
                                     upsample_rates=[4, 4, 2], upsample_kernel_sizes=[4, 8, 4],
                                     cond_d_vector_in_each_upsampling_layer=False),
    }
    for n, (tag, h) in enumerate(variants.items()):
        seed = 24 + n
        sd = O.make_state_dict(h, seed=seed)
        m = ref_v1.BigVGAN(_H(h), use_cuda_kernel=False)
        m.remove_weight_norm()
        missing, unexpected = m.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()}, strict=False)
        assert not unexpected and all(k.startswith("speaker_encoder.") for k in missing), (missing[:5], unexpected[:5])
        m.eval()
        emb = {}
        hk = m.speaker_encoder.register_forward_hook(lambda _m, _i, o: emb.__setitem__("e", o.detach().numpy().copy()))
        x = rng.standard_normal((2, 7 + n, h["gpt_dim"])).astype(np.float32)
        mel_ref = rng.standard_normal((2, 60, h["num_mels"])).astype(np.float32)
        with torch.no_grad():
            wav, loss = m(torch.from_numpy(x), torch.from_numpy(mel_ref))
        hk.remove()
        assert loss is None
        out[f"{tag}_x"], out[f"{tag}_emb"], out[f"{tag}_wav"], out[f"{tag}_seed"] = x, emb["e"][:, 0, :], wav.numpy(), np.array(seed)
        out[f"{tag}_h"] = np.array(json.dumps(h))
        print(tag, "embedding", emb["e"].shape, "wav", wav.shape, "rms", float(wav.pow(2).mean().sqrt()))
    # state-dict layout of the reference v1 module (weight-norm form), for the key/shape parity test
    m = ref_v1.BigVGAN(_H(dict(V1_MINI_H)), use_cuda_kernel=False)
    out["v1_state_dict_layout"] = np.array(json.dumps({k: list(v.shape) for k, v in m.state_dict().items()}))

    # ECAPA-TDNN speaker encoder (indextts/BigVGAN/ECAPA_TDNN.py): OUR module is built under a torch seed, its state dict
    # loaded into the reference's with strict=True, and the reference's embeddings stored (tests rebuild from the seed)
    from index_tts_b200 import ecapa as our_ecapa
    from indextts.BigVGAN.ECAPA_TDNN import ECAPA_TDNN as RefEcapa
    ours = build_seeded_ecapa(our_ecapa)
    ref = RefEcapa(100, lin_neurons=48)
    ref.load_state_dict(ours.state_dict(), strict=True)
    ref.eval()
    mel_ref = rng.standard_normal((3, 57, 100)).astype(np.float32)
    lens = np.array([1.0, 0.6, 0.85], dtype=np.float32)
    with torch.no_grad():
        e_full = ref(torch.from_numpy(mel_ref)).numpy()
        e_lens = ref(torch.from_numpy(mel_ref), torch.from_numpy(lens)).numpy()
    np.savez(os.path.join(HERE, "ecapa.npz"), mel_ref=mel_ref, lens=lens, emb=e_full, emb_lens=e_lens)
    print("ecapa embedding", e_full.shape, float(np.abs(e_full).mean()))
    np.savez(os.path.join(HERE, "variants.npz"), **out)
    print("written", os.path.join(HERE, "variants.npz"))


if __name__ == "__main__":
    main()
