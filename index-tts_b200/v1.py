"""The v1 (IndexTTS / IndexTTS-1.5) BigVGAN generator with the reference's Python surface and the B200-native forward.

Mirrors ``indextts/BigVGAN/models.py`` of the reference: ``AMPBlock1`` (:20-76), ``AMPBlock2`` (:79-127), ``BigVGAN``
(:130-258) -- same constructor arguments, module tree and state-dict keys (``conv_pre``, ``ups``, ``resblocks``,
``activation_post``, ``conv_post``, ``speaker_encoder``, ``cond_layer``, ``conds``), ``forward(x, mel_ref, lens)`` returning
``(wav, contrastive_loss)``, ``remove_weight_norm()``; it is what ``indextts/infer.py:29,125-132,489,659`` drives.

Against the v2 generator (``bigvgan.py``) the decoder differs in three places, all handled by ``bvg_forward_cond``
(include/bvg.h): the input is the channels-last GPT latent ``[B, T, gpt_dim]`` (models.py:221), a per-utterance vector
``cond_layer(speaker_embedding)`` is added after ``conv_pre`` (:236) and ``conds[i](speaker_embedding)`` after every
transposed conv (:244-245), and the output is always ``tanh`` (:256).  The 1x1 conditioning convs run as
``bvg_cond_linear_fwd``.  The ECAPA-TDNN speaker encoder (:202, once per call on the short reference mel, not part of the
vocoder's kernel path) is the torch module in ``ecapa.py``.
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch.nn.utils import remove_weight_norm, weight_norm

from . import _lib, activations, ops
from .bigvgan import (AMPBlock1, AMPBlock2 as _AMPBlock2, Conv1d, ConvTranspose1d, _GeneratorPlanMixin,
                      _select_activation1d, _vocoder_forward_cond)
from .ecapa import ECAPA_TDNN
from .utils import init_weights

LRELU_SLOPE = 0.1

# Shape class of the published v1 vocoder (24 kHz, 1024x up-sampling of the 1024-wide GPT latent, 512-wide speaker
# embedding, 100-band reference mel).  The v1 checkpoint's config file is a download, NOT part of the reference tree, so
# these values are not pinned by it; `BigVGAN(h)` takes whatever `cfg.bigvgan` holds (infer.py:125).  Used by bench.py.
DEFAULT_HPARAMS = {
    "resblock": "1",
    "upsample_rates": [4, 4, 4, 4, 2, 2],
    "upsample_kernel_sizes": [8, 8, 4, 4, 4, 4],
    "upsample_initial_channel": 1536,
    "resblock_kernel_sizes": [3, 7, 11],
    "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
    "activation": "snakebeta",
    "snake_logscale": True,
    "feat_upsample": False,
    "cond_d_vector_in_each_upsampling_layer": True,
    "speaker_embedding_dim": 512,
    "gpt_dim": 1024,
    "num_mels": 100,
    "sampling_rate": 24000,
}

__all__ = ["AMPBlock1", "AMPBlock2", "BigVGAN", "LRELU_SLOPE"]


class AMPBlock2(_AMPBlock2):
    """models.py:79-127: exactly TWO convs (``dilation[0]``, ``dilation[1]``), whatever the tuple's length."""

    def __init__(self, h, channels, kernel_size=3, dilation=(1, 3), activation=None):
        super().__init__(h, channels, kernel_size, tuple(dilation)[:2], activation)


class BigVGAN(_GeneratorPlanMixin, torch.nn.Module):
    """Reference ``indextts/BigVGAN/models.py:130-258``.

    Args:
        h: hyper-parameters with attribute access (``cfg.bigvgan`` in ``infer.py:125``); mutated like the reference:
            ``h["use_cuda_kernel"]``.
        use_cuda_kernel (bool): accepted and stored; both values run the sm_100a kernels.
    """

    def __init__(self, h, use_cuda_kernel=False):
        super().__init__()
        self.h = h
        self.h["use_cuda_kernel"] = use_cuda_kernel

        self.num_kernels = len(h.resblock_kernel_sizes)
        self.num_upsamples = len(h.upsample_rates)
        self.feat_upsample = h.feat_upsample
        self.cond_in_each_up_layer = h.cond_d_vector_in_each_upsampling_layer

        self.conv_pre = weight_norm(Conv1d(h.gpt_dim, h.upsample_initial_channel, 7, 1, padding=3))
        resblock = AMPBlock1 if h.resblock == "1" else AMPBlock2  # models.py:156: anything else is AMPBlock2

        self.ups = nn.ModuleList()
        for i, (u, k) in enumerate(zip(h.upsample_rates, h.upsample_kernel_sizes)):
            self.ups.append(nn.ModuleList([
                weight_norm(ConvTranspose1d(h.upsample_initial_channel // (2 ** i),
                                            h.upsample_initial_channel // (2 ** (i + 1)), k, u, padding=(k - u) // 2))]))

        self.resblocks = nn.ModuleList()
        for i in range(len(self.ups)):
            ch = h.upsample_initial_channel // (2 ** (i + 1))
            for j, (k, d) in enumerate(zip(h.resblock_kernel_sizes, h.resblock_dilation_sizes)):
                self.resblocks.append(resblock(self.h, ch, k, d, activation=h.activation))
        Activation1d = _select_activation1d({"use_cuda_kernel": use_cuda_kernel})

        if h.activation == "snake":
            self.activation_post = Activation1d(activation=activations.Snake(ch, alpha_logscale=h.snake_logscale))
        elif h.activation == "snakebeta":
            self.activation_post = Activation1d(activation=activations.SnakeBeta(ch, alpha_logscale=h.snake_logscale))
        else:
            raise NotImplementedError(
                "activation incorrectly specified. check the config file and look for 'activation'.")

        self.conv_post = weight_norm(Conv1d(ch, 1, 7, 1, padding=3))

        for i in range(len(self.ups)):
            self.ups[i].apply(init_weights)
        self.conv_post.apply(init_weights)

        self.speaker_encoder = ECAPA_TDNN(h.num_mels, lin_neurons=h.speaker_embedding_dim)
        self.cond_layer = nn.Conv1d(h.speaker_embedding_dim, h.upsample_initial_channel, 1)
        if self.cond_in_each_up_layer:
            self.conds = nn.ModuleList()
            for i in range(len(self.ups)):
                ch = h.upsample_initial_channel // (2 ** (i + 1))
                self.conds.append(nn.Conv1d(h.speaker_embedding_dim, ch, 1))

        self.use_tanh_at_final = True  # models.py:256
        self._init_plan_state()

    # ---- forward ---------------------------------------------------------------------------------------------
    def forward_with_embedding(self, x, speaker_embedding):
        """The decoder half of ``forward`` (models.py:212-258): ``x`` ``[B, T, gpt_dim]`` CUDA tensor,
        ``speaker_embedding`` ``[B, D]`` (or ``[B, 1, D]``) -> ``[B, 1, T * prod(upsample_rates)]`` fp32 in [-1, 1]."""
        if x.dim() != 3 or x.shape[2] != self.conv_pre.in_channels:
            raise ValueError(f"expected a latent of shape [B, T, {self.conv_pre.in_channels}], got {tuple(x.shape)}")
        if not x.is_cuda:
            raise _lib.BvgError("BigVGAN.forward: libbvg has no CPU path; move the module and the inputs to a B200 "
                                f"(got device {x.device})")
        _lib.require_device(x.device.index if x.device.index is not None else torch.cuda.current_device())
        e = speaker_embedding.reshape(speaker_embedding.shape[0], self.cond_layer.in_channels).float()
        if e.shape[0] != x.shape[0]:
            raise ValueError(f"{e.shape[0]} speaker embeddings for a batch of {x.shape[0]}")
        if self.feat_upsample:
            # models.py:214-219: 4x linear interpolation of the latent along time (off in the published config); a host-side
            # torch op on the tiny latent, before the kernels
            x = torch.nn.functional.interpolate(x.transpose(1, 2), scale_factor=[4], mode="linear").transpose(1, 2)
        x = x.float().contiguous()
        plan = self._plan(x.device)
        return _vocoder_forward_cond(x, self._cond_vectors(e, x), plan.plan_id, plan.hop)

    def _cond_vectors(self, e, like):
        conds = [ops.cond_linear(e, self.cond_layer.weight, self.cond_layer.bias)]
        for i in range(self.num_upsamples):
            conds.append(ops.cond_linear(e, self.conds[i].weight, self.conds[i].bias) if self.cond_in_each_up_layer
                         else like.new_empty(0))
        return conds

    def profile_with_embedding(self, x, speaker_embedding):
        """Measurement helper (bench.py): ``forward_with_embedding`` with a CUDA event pair around every launch."""
        from .bigvgan import profile_forward_cond
        x = x.float().contiguous()
        e = speaker_embedding.reshape(speaker_embedding.shape[0], self.cond_layer.in_channels).float()
        return profile_forward_cond(self._plan(x.device), x, self._cond_vectors(e, x))

    def forward(self, x, mel_ref, lens=None):
        """``x`` ``[B, T, gpt_dim]``, ``mel_ref`` ``[B, T_ref, num_mels]`` -> ``(wav [B, 1, T * hop], None)``
        (models.py:201-258)."""
        speaker_embedding = self.speaker_encoder(mel_ref, lens)
        n_batch = x.size(0)
        contrastive_loss = None
        if n_batch * 2 == speaker_embedding.size(0):
            # models.py:205-209 reads self.logit_scale, which the reference commented out (:199): same failure here
            raise AttributeError("'BigVGAN' object has no attribute 'logit_scale'")
        return self.forward_with_embedding(x, speaker_embedding), contrastive_loss

    def remove_weight_norm(self):
        self._bump_generation()
        print('Removing weight norm...')
        for l in self.ups:
            for l_i in l:
                remove_weight_norm(l_i)
        for l in self.resblocks:
            l.remove_weight_norm()
        remove_weight_norm(self.conv_pre)
        remove_weight_norm(self.conv_post)
