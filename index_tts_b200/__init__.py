"""Importable alias of the ``index-tts_b200/`` package directory (a hyphen is not a valid module name).

``import index_tts_b200.bigvgan`` resolves to ``index-tts_b200/bigvgan.py``.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "index-tts_b200")
if not _os.path.isdir(_real):  # pragma: no cover
    raise ImportError(f"package directory {_real} is missing")
__path__ = [_real]
__version__ = "0.1.0"
