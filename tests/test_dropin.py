"""the unmodified reference call sites (infer_v2.py:42, 174-182, 259-264, 856-866) against this package
through `index_tts_b200.dropin.install()`.  CPU part: module resolution; GPU part: the vocoder tail of
`IndexTTS2.infer_generator` executed statement for statement and checked against the oracle."""
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

from helpers import MINI_H, O, ROOT, build_model, snr_db

_RESOLVE = textwrap.dedent(
    """
    import sys
    sys.path.insert(0, {root!r})
    {extra}
    import index_tts_b200.dropin as dropin
    dropin.install()
    # infer_v2.py:42
    from indextts.s2mel.modules.bigvgan import bigvgan
    # infer_v2.py:176
    from indextts.s2mel.modules.bigvgan.alias_free_activation.cuda import activation1d
    import index_tts_b200.bigvgan as ours
    assert bigvgan is ours, bigvgan
    assert hasattr(activation1d, "anti_alias_activation_cuda") and hasattr(activation1d, "Activation1d")
    from indextts.s2mel.modules.bigvgan.alias_free_activation.torch.act import Activation1d
    from indextts.s2mel.modules.bigvgan.activations import Snake, SnakeBeta
    from indextts.s2mel.modules.bigvgan.env import AttrDict
    from indextts.s2mel.modules.bigvgan.utils import get_padding, init_weights
    assert get_padding(11, 5) == 25
    # infer_v2.py:259-264 with a local checkpoint directory
    m = bigvgan.BigVGAN.from_pretrained({ckpt!r}, use_cuda_kernel=True)
    m.remove_weight_norm()
    m.eval()
    assert m.h["use_cuda_kernel"] is True
    # v1 pipeline: infer.py:29 and the preload at infer.py:116-121
    from indextts.BigVGAN.models import BigVGAN as Generator
    import index_tts_b200.v1 as ours_v1
    assert Generator is ours_v1.BigVGAN
    from indextts.BigVGAN.alias_free_activation.cuda import load
    assert "libbvg" in repr(load.load())
    dropin.uninstall()
    assert "indextts.s2mel.modules.bigvgan.bigvgan" not in sys.modules and "indextts.BigVGAN.models" not in sys.modules
    print("ok")
    """)


def _make_ckpt(tmp_path):
    from index_tts_b200 import bigvgan as B
    torch.manual_seed(5)
    m = B.BigVGAN(B.AttrDict(dict(MINI_H)), use_cuda_kernel=False)
    m._save_pretrained(tmp_path)
    return str(tmp_path)


@pytest.mark.parametrize("with_reference_tree", [False, True])
def test_reference_import_lines_resolve_to_this_package(tmp_path, with_reference_tree):
    if with_reference_tree and not os.path.isdir("/root/reference/indextts"):
        pytest.skip("reference tree not present on this box")
    extra = "sys.path.insert(0, '/root/reference')" if with_reference_tree else ""
    code = _RESOLVE.format(root=ROOT, extra=extra, ckpt=_make_ckpt(tmp_path))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.strip().endswith("ok")


@pytest.mark.gpu
def test_infer_v2_vocoder_tail_unchanged():
    """infer_v2.py:856-866 executed as written (time-sliced non-contiguous mel, .float(), squeeze/unsqueeze, the
    32767 clamp) with `self.bigvgan` built through the reference's module path."""
    import index_tts_b200.dropin as dropin
    dropin.install()
    try:
        from indextts.s2mel.modules.bigvgan import bigvgan  # noqa: F401  (infer_v2.py:42)
        sd = O.make_state_dict(MINI_H, seed=21)
        self_bigvgan = build_model(MINI_H, sd, precision="fp32")
        assert isinstance(self_bigvgan, bigvgan.BigVGAN)
        rng = np.random.default_rng(3)
        ref_len = 17
        vc_full = torch.from_numpy((4.0 * rng.standard_normal((1, 80, ref_len + 29))).astype(np.float32)).cuda()
        with torch.no_grad():
            vc_target = vc_full[:, :, ref_len:]                              # infer_v2.py:856
            assert not vc_target.is_contiguous()
            wav = self_bigvgan(vc_target.float()).squeeze().unsqueeze(0)     # infer_v2.py:860
            wav = wav.squeeze(1)                                             # infer_v2.py:864
            wav = torch.clamp(32767 * wav, -32767.0, 32767.0)                # infer_v2.py:866
        ref = O.forward(sd, vc_full[:, :, ref_len:].cpu().numpy(), MINI_H, np.float64)
        ref = np.clip(32767.0 * ref.reshape(1, -1), -32767.0, 32767.0)
        got = wav.cpu().numpy()
        assert got.shape == ref.shape == (1, 29 * 256)
        assert float(np.abs(got - ref).max()) <= 32767 * 1e-4
        assert snr_db(ref, got) >= 80.0
        pcm = wav.type(torch.int16)                                          # infer_v2.py:912
        assert int((pcm.cpu().numpy().astype(np.int64) - np.round(ref).astype(np.int64)).__abs__().max()) <= 4
        # the same conversion fused into the tail kernel -- identical bits, non-contiguous mel included
        with torch.no_grad():
            fused = self_bigvgan.forward_pcm16(vc_target.float()).squeeze().unsqueeze(0)
        assert fused.dtype == torch.int16 and torch.equal(fused, pcm)
    finally:
        dropin.uninstall()


@pytest.mark.slow
def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` (the driver's CPU arm) needs no GPU: exactly one JSON line on stdout with the contract's
    keys, the CPU implementation named in `cpu_baseline.kind` and the bounded sample it really ran."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, BVG_REF_BUDGET_S="4")  # halves the sample down to ~108 frames: seconds instead of minutes
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=900, env=env, cwd=root)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[-1000:]
    j = json.loads(lines[0])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "config",
                "cpu_baseline", "e2e"):
        assert key in j, key
    assert j["impl"] == "reference" and j["unit"] == "audio-s/s" and j["value"] > 0
    assert j["cpu_baseline"]["kind"] in ("reference", "port") and j["cpu_baseline"]["cores"] >= 1
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["e2e"]["d2h_bytes_per_step"] == 0
