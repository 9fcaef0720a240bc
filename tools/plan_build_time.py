"""Developer tool: how long does folding weight-norm + packing every weight of bigvgan_v2_22khz_80band_256x take?
(SURVEY.md 8(f) N3 asks whether an on-disk cache of packed weights would pay at start-up.)"""
import os, sys, time, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
with contextlib.redirect_stdout(sys.stderr):
    model = bench.make_state_dict_torch()
t0 = time.perf_counter(); model = model.to("cuda"); torch.cuda.synchronize(); t1 = time.perf_counter()
print(f"module.to(cuda) (449 MB fp32): {1e3 * (t1 - t0):.1f} ms")
for prec in ("bf16", "fp32", "bf16"):
    model.set_precision(prec)
    model.__dict__["_bvg_plan"] = None; model.__dict__["_bvg_plan_key"] = None
    for m in model.modules():
        if "_bvg_cache" in m.__dict__:
            m.__dict__["_bvg_cache"].clear()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    model._plan(torch.device("cuda", 0)); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"plan build ({prec}): {1e3 * (t1 - t0):.1f} ms")
mel = torch.randn(1, 80, 431, device="cuda")
model.set_precision("bf16")
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.no_grad(): model(mel)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    print(f"forward cfg1 call {i}: {1e3 * (t1 - t0):.2f} ms")
