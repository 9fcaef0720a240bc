"""No-GPU checks of the drop-in boundary: libbvg.so loads, exports every symbol include/bvg.h declares, the ctypes
model descriptor matches the C struct, and host-side argument validation works without a device."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "bvg.h")).read()
    return sorted(set(re.findall(r"BVG_API\s+[\w\s\*]+?\b(bvg_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from index_tts_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/bvg.h but not exported by libbvg.so"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes signature"
    assert lib.bvg_version() == 100


def test_model_desc_layout_matches_c_struct(tmp_path):
    """Compile a tiny C program against include/bvg.h and compare sizeof/offsetof with the ctypes mirror."""
    from index_tts_b200 import _lib
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "bvg.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(bvg_model_desc), offsetof(bvg_model_desc, conv_pre), offsetof(bvg_model_desc, convs1),'
                   'offsetof(bvg_model_desc, acts), offsetof(bvg_model_desc, conv_post), sizeof(bvg_conv_desc));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    M = _lib.ModelDesc
    assert got == [ctypes.sizeof(M), M.conv_pre.offset, M.convs1.offset, M.acts.offset, M.conv_post.offset,
                   ctypes.sizeof(_lib.ConvDesc)]


def test_argument_validation_without_gpu():
    from index_tts_b200 import _lib
    lib = _lib.load()
    assert lib.bvg_activation1d_fwd(None, None, None, None, 1, 1, 1, 1, 1, 0, 1, None) == -1
    assert b"null" in lib.bvg_last_error()
    buf = (ctypes.c_float * 64)()
    p = ctypes.addressof(buf)
    assert lib.bvg_activation1d_fwd(p, p, p, p, 1, 1, 8, 8, 8, 7, 1, None) == -2      # bad dtype
    assert lib.bvg_activation1d_fwd(p, p, p, p, 0, 4, 8, 32, 8, 0, 1, None) == 0      # empty: no-op, no launch
    assert lib.bvg_pack_convtranspose1d_weight(p, p, 8, 8, 7, 2, 0, None) == -2       # K % u != 0
    assert lib.bvg_workspace_bytes(ctypes.byref(_lib.ModelDesc()), 1, 10) == 0        # empty descriptor is invalid


def test_module_surface_and_cpu_forward_raises():
    from index_tts_b200 import bigvgan as B, _lib
    from index_tts_b200.alias_free_activation.cuda import activation1d
    from helpers import MINI_H
    assert hasattr(activation1d, "anti_alias_activation_cuda") and hasattr(activation1d, "Activation1d")
    h = B.AttrDict(dict(MINI_H))
    m = B.BigVGAN(h, use_cuda_kernel=True)
    assert h["use_cuda_kernel"] is True
    keys = list(m.state_dict().keys())
    assert len(keys) == 783 and "conv_pre.weight_g" in keys and "resblocks.17.activations.5.downsample.lowpass.filter" in keys
    m.remove_weight_norm()
    m.remove_weight_norm()  # idempotent ("already removed" message), reference bigvgan.py:398-400
    assert len(m.state_dict()) == 667
    with pytest.raises(_lib.BvgError):
        m(torch.zeros(1, 80, 4))
    with pytest.raises(ValueError):
        B.BigVGAN(B.AttrDict(dict(MINI_H, resblock="3")))
    with pytest.raises(NotImplementedError):
        B.BigVGAN(B.AttrDict(dict(MINI_H, activation="relu")))
    assert m.resblocks[0].activations[0].upsample.filter.shape == (1, 1, 12)


def test_save_and_load_local_directory(tmp_path):
    from index_tts_b200 import bigvgan as B
    from helpers import MINI_H
    torch.manual_seed(3)
    m = B.BigVGAN(B.AttrDict(dict(MINI_H)))
    m._save_pretrained(tmp_path)
    m2 = B.BigVGAN._from_pretrained(model_id=str(tmp_path), use_cuda_kernel=False)
    for (k1, v1), (k2, v2) in zip(m.state_dict().items(), m2.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2)
    # checkpoint without weight-norm keys: load retries after remove_weight_norm() (reference bigvgan.py:483-490)
    m.remove_weight_norm()
    m._save_pretrained(tmp_path)
    m3 = B.BigVGAN._from_pretrained(model_id=str(tmp_path))
    assert len(m3.state_dict()) == 667


@pytest.mark.skipif(not os.path.isdir("/root/reference/indextts"), reason="reference tree not present on this box")
def test_seeded_construction_matches_reference_state_dict():
    """Same RNG consumption as the reference's constructor: torch.manual_seed(s); BigVGAN(h) gives identical weights."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden as G
    from index_tts_b200 import bigvgan as B
    from helpers import MINI_H
    rb, _, _ = G.import_reference()
    torch.manual_seed(1234)
    ours = B.BigVGAN(B.AttrDict(dict(MINI_H))).state_dict()
    torch.manual_seed(1234)
    ref = rb.BigVGAN(rb.AttrDict(dict(MINI_H))).state_dict()
    assert list(ours.keys()) == list(ref.keys())
    assert all(torch.equal(ours[k], ref[k]) for k in ref)
