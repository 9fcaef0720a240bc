"""The reference's operator sequence on the SAME B200 through stock PyTorch (cuDNN / ATen CUDA kernels), timed next to
libbvg ("same-box GPU baselines to list beside ours").  The reference itself cannot travel to the GPU
box, so this uses oracle/torch_port.py -- the reference's forward restated op for op on torch -- moved to the GPU.
It is context for the roofline discussion, not a parity oracle (TF32 / bf16 eager are not fp32-exact): the only
assertions are that both paths agree loosely and that the hand-written path is faster.  Results are printed and, when
`gpurun_out/` exists, written to gpurun_out/torch_gpu_baseline.json."""
import json
import os
import time

import numpy as np
import pytest
import torch

from helpers import O, ROOT, build_model, snr_db
from oracle import torch_port as TP

pytestmark = pytest.mark.gpu


def _time(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def test_stock_pytorch_on_the_same_gpu():
    sd_np = O.make_state_dict(O.DEFAULT_HPARAMS, seed=1)
    ours = build_model(O.DEFAULT_HPARAMS, sd_np, precision="bf16")
    sd = {k: v.cuda() for k, v in TP.prepare(sd_np).items()}
    sd16 = {k: (v.bfloat16() if v.is_floating_point() else v) for k, v in sd.items()}
    rows = []
    for name, B, T, reps in (("cfg1: 1 x 5 s", 1, 431, 5), ("cfg2/4: 4 x 10 s", 4, 862, 2)):
        mel = torch.randn(B, 80, T, generator=torch.Generator().manual_seed(B)).cuda()
        audio_s = B * T * 256 / 22050.0
        res = {"workload": name, "audio_s": audio_s}
        with torch.no_grad():
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False
            ref32 = TP.forward(sd, mel)
            res["torch_fp32_ms"] = _time(lambda: TP.forward(sd, mel), reps)
            torch.backends.cudnn.allow_tf32 = True
            res["torch_tf32_ms"] = _time(lambda: TP.forward(sd, mel), reps)
            res["torch_bf16_ms"] = _time(lambda: TP.forward(sd16, mel.bfloat16()), reps)
            got = ours(mel)
            res["libbvg_bf16_ms"] = _time(lambda: ours(mel), 10)
        res["libbvg_bf16_snr_db_vs_torch_fp32"] = snr_db(ref32.float().cpu().numpy(), got.cpu().numpy())
        for k in ("torch_fp32_ms", "torch_tf32_ms", "torch_bf16_ms", "libbvg_bf16_ms"):
            res[k.replace("_ms", "_audio_s_per_s")] = audio_s / (res[k] * 1e-3)
        print(json.dumps(res))
        rows.append(res)
        assert res["libbvg_bf16_snr_db_vs_torch_fp32"] >= 40.0
        assert res["libbvg_bf16_ms"] < min(res["torch_fp32_ms"], res["torch_tf32_ms"], res["torch_bf16_ms"])
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        json.dump({"when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "rows": rows},
                  open(os.path.join(out_dir, "torch_gpu_baseline.json"), "w"), indent=1)
