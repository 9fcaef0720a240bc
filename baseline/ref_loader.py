"""Loader for the UNMODIFIED reference vocoder installed by ``baseline/install_ref.sh`` into ``baseline/_ref/``.

Used only by ``bench.py --impl reference`` / its ``cpu_baseline`` leg, ``tools/`` and ``tests/`` -- never by the product
package.  The reference package imports matplotlib (``utils.py``) and librosa (``meldataset.py``) at module scope for
plotting / mel extraction helpers the vocoder forward never calls; neither is in this image, so import-only stubs stand in
for them.  The reference's own files are byte-for-byte copies.
"""
import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")
K1_SO = os.path.join(REF_ROOT, "k1", "anti_alias_activation_cuda.so")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "indextts", "s2mel", "modules", "bigvgan", "bigvgan.py"))


def import_reference():
    """Returns the reference's ``(bigvgan, activations, Activation1d)`` modules/classes from ``baseline/_ref``."""
    if not available():
        raise ImportError("baseline/_ref is empty: run baseline/install_ref.sh where /root/reference exists")
    for name in ["matplotlib", "matplotlib.pylab", "librosa", "librosa.util", "librosa.filters"]:
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:
                sys.modules[name] = types.ModuleType(name)
    mpl = sys.modules["matplotlib"]
    if not hasattr(mpl, "use"):
        mpl.use = lambda *a, **k: None
    if not hasattr(sys.modules["librosa.util"], "normalize"):
        sys.modules["librosa.util"].normalize = None
    if not hasattr(sys.modules["librosa.filters"], "mel"):
        sys.modules["librosa.filters"].mel = None
    # a drop-in alias (index_tts_b200.dropin.install()) must not shadow the real reference here
    for k in [k for k in sys.modules if k == "indextts" or k.startswith("indextts.")]:
        mod = sys.modules[k]
        if not (getattr(mod, "__file__", None) or "").startswith(REF_ROOT) and \
                not any(str(p).startswith(REF_ROOT) for p in getattr(mod, "__path__", [])):
            raise ImportError(f"{k} is already imported from elsewhere ({getattr(mod, '__file__', None)})")
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from indextts.s2mel.modules.bigvgan import activations, bigvgan
    from indextts.s2mel.modules.bigvgan.alias_free_activation.torch.act import Activation1d
    return bigvgan, activations, Activation1d


def load_k1():
    """The reference's fused extension (``anti_alias_activation_cuda``: ``forward(x, up_ftr, down_ftr, alpha, beta)``) as
    built for sm_100 by ``install_ref.sh``, or ``None`` with the reason."""
    if not os.path.isfile(K1_SO):
        return None, f"{K1_SO} not built"
    try:
        import torch  # noqa: F401  (the extension links against libtorch)
        spec = importlib.util.spec_from_file_location("anti_alias_activation_cuda", K1_SO)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod, None
    except Exception as e:  # pragma: no cover
        return None, f"{type(e).__name__}: {e}"
