"""Developer tool: per-layer CUDA-event times of the cfg2 forward (bvg_forward_profile), min over a few passes, aggregated
by (kind, stage, k, residual?).  Run once per environment-knob setting to A/B kernels inside the real plan.

    BVG_CONV_GROUPS=1 python tools/layer_ab.py [passes]
"""
import os, sys, collections, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from index_tts_b200 import bigvgan as B
passes = int(sys.argv[1]) if len(sys.argv) > 1 else 4
with contextlib.redirect_stdout(sys.stderr):
    model = bench.make_state_dict_torch().to("cuda")
model.set_precision("bf16")
mel = torch.randn(16, 80, 862, generator=torch.Generator().manual_seed(1)).cuda()
best = None
for _ in range(passes):
    _, recs = B.profile_forward(model, mel)
    ms = [r["ms"] for r in recs]
    best = ms if best is None else [min(a, b) for a, b in zip(best, ms)]
agg = collections.OrderedDict()
seen = collections.Counter()
for r, ms in zip(recs, best):
    key = (r["kind"][:6], r["stage"], r["k"])
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1; a[1] += ms
tot = collections.Counter()
for (kind, stage, k), (n, ms) in agg.items():
    tot[kind] += ms
    print(f"{kind:7s} st{stage:2d} k{k:2d} n={n:2d} {ms:7.3f} ms  ({ms / n * 1e3:6.1f} us each)")
if os.environ.get("LAYER_AB_DETAIL"):
    # every launch against its own roofline: max(flops / sustained bf16 peak, bytes / HBM)
    for r, ms in zip(recs, best):
        floor = max(r["flops"] / 1396.8e12, r["bytes"] / 6555.2e9) * 1e3
        print(f"  {r['kind'][:6]:7s} st{r['stage']:2d} C{r['cin']:4d} k{r['k']:2d} d{r['dilation']} {ms * 1e3:7.1f} us  floor {floor * 1e3:6.1f} us"
              f"  frac {floor / ms if ms else 0:4.2f}  lost {1e3 * (ms - floor):6.1f} us")
print("totals:", {k: round(v, 3) for k, v in tot.items()}, "sum", round(sum(tot.values()), 3))
