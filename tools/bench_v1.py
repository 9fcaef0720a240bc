#!/usr/bin/env python
"""v1 generator timing (decoder alone, embedding given): python tools/bench_v1.py [--reps N] [--precision bf16]
Prints one JSON line; run twice with BVG_COND_FUSED=0/1 for the conditioning-fusion A/B."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from index_tts_b200 import bigvgan as B, ops, v1 as V1  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=8)
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--frames", type=int, default=235)
ap.add_argument("--precision", default="bf16")
a = ap.parse_args()
dev = torch.device("cuda:0")
h = B.AttrDict(dict(V1.DEFAULT_HPARAMS))
torch.manual_seed(4321)
m = V1.BigVGAN(h, use_cuda_kernel=True)
m.remove_weight_norm()
m = m.to(dev).eval().set_precision(a.precision)
g = torch.Generator().manual_seed(3)
x = torch.randn(a.batch, a.frames, h.gpt_dim, generator=g).to(dev)
e = torch.randn(a.batch, h.speaker_embedding_dim, generator=g).to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
with torch.no_grad():
    for _ in range(3):
        m.forward_with_embedding(x, e)
    ops.reset_launch_count()
    m.forward_with_embedding(x, e)
    launches = ops.launch_count()
    ms = []
    for _ in range(a.reps):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        m.forward_with_embedding(x, e)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
ms.sort()
audio = a.batch * a.frames * 1024 / 24000.0
print(json.dumps({"cond_fused": os.environ.get("BVG_COND_FUSED", "1"), "precision": a.precision, "launches": launches,
                  "ms_median": ms[len(ms) // 2], "ms_min": ms[0], "audio_s_per_s": audio / (ms[len(ms) // 2] * 1e-3)}))
