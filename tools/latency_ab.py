#!/usr/bin/env python
"""Single-utterance latency A/B (developer tool): python tools/latency_ab.py [frames ...]; run under different env knobs."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from index_tts_b200 import bigvgan as B  # noqa: E402

frames = [int(a) for a in sys.argv[1:]] or [431]
dev = torch.device("cuda:0")
h = B.load_hparams_from_json(os.path.join(os.path.dirname(B.__file__), "config.json"))
torch.manual_seed(1234)
m = B.BigVGAN(h, use_cuda_kernel=True)
m.remove_weight_norm()
m = m.to(dev).eval().set_precision("bf16")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for T in frames:
    mel = torch.randn(1, 80, T, generator=torch.Generator().manual_seed(0)).to(dev)
    res = {}
    for mode in ("eager", "graph"):
        m.set_cuda_graphs(8192 if mode == "graph" else 0)
        with torch.no_grad():
            for _ in range(5):
                m(mel)
            ms = []
            for _ in range(20):
                flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); m(mel); e1.record()
                torch.cuda.synchronize()
                ms.append(e0.elapsed_time(e1))
        ms.sort()
        res[mode] = ms[len(ms) // 2]
    print(f"T={T} pair_smallm={os.environ.get('BVG_CONV_PAIR_SMALLM', '1')} eager {res['eager']:.3f} ms graph {res['graph']:.3f} ms")
