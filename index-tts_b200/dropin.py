"""Zero-line-change swap-in for the reference.

``install()`` registers this package under the reference's own module path, so that the unmodified
``indextts/infer_v2.py`` -- ``from indextts.s2mel.modules.bigvgan import bigvgan`` (infer_v2.py:42), the kernel preload
``from indextts.s2mel.modules.bigvgan.alias_free_activation.cuda import activation1d`` (infer_v2.py:174-182), and
``bigvgan.BigVGAN.from_pretrained(...).to(device)``, ``remove_weight_norm()``, ``eval()``, ``self.bigvgan(mel)``
(infer_v2.py:259-264, 860) -- resolves to the B200 implementation::

    import index_tts_b200.dropin as dropin
    dropin.install()            # before `import indextts.infer_v2`
    from indextts.infer_v2 import IndexTTS2

Nothing is copied or patched inside the reference tree: the aliases live in ``sys.modules`` only.  ``uninstall()``
removes them again.  Parent packages that are not importable (the reference tree is absent, as on the GPU box) are
created as empty namespace stubs, so the same two import lines work with or without ``/root/reference`` on ``sys.path``.
"""
import importlib
import sys
import types

REFERENCE_PACKAGE = "indextts.s2mel.modules.bigvgan"

# reference module (relative to REFERENCE_PACKAGE) -> module of this package
_ALIASES = {
    "": "index_tts_b200",
    ".bigvgan": "index_tts_b200.bigvgan",
    ".activations": "index_tts_b200.activations",
    ".env": "index_tts_b200.env",
    ".utils": "index_tts_b200.utils",
    ".alias_free_activation": "index_tts_b200.alias_free_activation",
    ".alias_free_activation.torch": "index_tts_b200.alias_free_activation.torch",
    ".alias_free_activation.torch.act": "index_tts_b200.alias_free_activation.torch.act",
    ".alias_free_activation.torch.filter": "index_tts_b200.alias_free_activation.torch.filter",
    ".alias_free_activation.torch.resample": "index_tts_b200.alias_free_activation.torch.resample",
    ".alias_free_activation.cuda": "index_tts_b200.alias_free_activation.cuda",
    ".alias_free_activation.cuda.activation1d": "index_tts_b200.alias_free_activation.cuda.activation1d",
    ".alias_free_activation.cuda.load": "index_tts_b200.alias_free_activation.cuda.load",
}
# the v1 pipeline's vocoder (indextts/infer.py:29 `from indextts.BigVGAN.models import BigVGAN as Generator`, :116-121 the
# kernel preload `from indextts.BigVGAN.alias_free_activation.cuda import load; load.load()`)
V1_REFERENCE_PACKAGE = "indextts.BigVGAN"
_V1_ALIASES = {
    "": "index_tts_b200",
    ".models": "index_tts_b200.v1",
    ".ECAPA_TDNN": "index_tts_b200.ecapa",
    ".activations": "index_tts_b200.activations",
    ".utils": "index_tts_b200.utils",
    ".alias_free_torch": "index_tts_b200.alias_free_activation.torch",
    ".alias_free_activation": "index_tts_b200.alias_free_activation",
    ".alias_free_activation.torch": "index_tts_b200.alias_free_activation.torch",
    ".alias_free_activation.cuda": "index_tts_b200.alias_free_activation.cuda",
    ".alias_free_activation.cuda.activation1d": "index_tts_b200.alias_free_activation.cuda.activation1d",
    ".alias_free_activation.cuda.load": "index_tts_b200.alias_free_activation.cuda.load",
}
_installed = {}


def _ensure_parent_packages():
    """Make ``indextts``, ``indextts.s2mel``, ``indextts.s2mel.modules`` importable parents: the real ones when the
    reference tree is on ``sys.path``, otherwise empty stubs."""
    parts = REFERENCE_PACKAGE.split(".")[:-1]
    for i in range(1, len(parts) + 1):
        name = ".".join(parts[:i])
        if name in sys.modules:
            continue
        try:
            importlib.import_module(name)
        except Exception:
            stub = types.ModuleType(name)
            stub.__path__ = []  # a package, so that `import a.b.c` walks through it
            sys.modules[name] = stub
            _installed[name] = stub
            if i > 1:
                setattr(sys.modules[".".join(parts[:i - 1])], parts[i - 1], stub)


def install(preload_kernels: bool = True, precision: str = None):
    """Alias this package as ``indextts.s2mel.modules.bigvgan`` and, for the v1 pipeline, ``indextts.BigVGAN`` (idempotent).  With ``preload_kernels`` the shared
    library is loaded right away, so a missing ``libbvg.so`` fails here, loudly, instead of at the first forward.

    ``precision`` ("bf16" | "fp32" | "fp32_simt") sets the arithmetic of every generator constructed afterwards; the
    reference runs this path in fp32 (infer_v2.py:830-831, 860), so the choice is logged either way: the default
    (``$BVG_PRECISION`` or "bf16") trades exact fp32 for the bf16 tensor-core mode (>= 40 dB against the fp32 result)."""
    import logging
    from . import bigvgan as _bv, ops as _ops
    if precision is not None:
        _ops.precision_code(precision)  # validates the name
        _bv.DEFAULT_PRECISION = precision
    logging.getLogger("index_tts_b200").info("dropin.install: vocoder precision mode %r (the reference runs fp32; "
                                             "set_precision() or install(precision=...) changes it)", _bv.DEFAULT_PRECISION)
    _ensure_parent_packages()
    table = [(REFERENCE_PACKAGE + rel, ours) for rel, ours in _ALIASES.items()]
    table += [(V1_REFERENCE_PACKAGE + rel, ours) for rel, ours in _V1_ALIASES.items()]
    for name, ours in table:
        mod = importlib.import_module(ours)
        sys.modules[name] = mod
        _installed[name] = mod
        parent, _, leaf = name.rpartition(".")
        if parent in sys.modules:
            try:
                setattr(sys.modules[parent], leaf, mod)
            except Exception:
                pass
    if preload_kernels:
        from . import _lib
        _lib.load()
    return sys.modules[REFERENCE_PACKAGE + ".bigvgan"]


def uninstall():
    for name, mod in list(_installed.items()):
        if sys.modules.get(name) is mod:
            del sys.modules[name]
    _installed.clear()
