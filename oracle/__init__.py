"""CPU oracle for the BigVGAN-v2 vocoder forward path (TEST INFRASTRUCTURE ONLY).

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker or the timed CPU baseline.
The product path (``index-tts_b200/``) never imports this package and raises when
its CUDA library is missing.

Parity status: the reference ships no golden vectors for this path (SURVEY.md §4,
§8c), so the oracle is pinned against outputs of the reference's own torch
implementation imported in the authoring container (``tests/golden/make_golden.py``
-> ``tests/golden/*.npz``).
"""
from .bigvgan_oracle import *  # noqa: F401,F403
