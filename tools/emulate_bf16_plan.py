"""CPU emulation of the BVG_BF16_TC numerics plan (developer tool; needs /root/reference).

bf16-rounded MMA operands (activation outputs and weights), exact products + fp32 accumulation, fp32 residual
stream, fp32 Activation1d arithmetic; optional bf16 rounding of the convs1 output (the tensor between c1 and a2).
Reports SNR against the reference's own fp32 forward so the rounding points can be chosen before the kernels exist.
"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden"))
import numpy as np, torch
import torch.nn.functional as F
import make_golden as G
from oracle import bigvgan_oracle as O

def r16(x):
    return x.to(torch.bfloat16).to(torch.float32)

def run(ref, mel, y1_bf16, pre_bf16=True, res16_from_stage=99, x16_from_stage=99):
    h = ref.h
    nk = ref.num_kernels
    with torch.no_grad():
        x = F.conv1d(r16(mel), r16(ref.conv_pre.weight), ref.conv_pre.bias, padding=3)
        for i in range(ref.num_upsamples):
            up = ref.ups[i][0]
            x = F.conv_transpose1d(r16(x), r16(up.weight), up.bias, stride=up.stride, padding=up.padding)
            xs = None
            for j in range(nk):
                rb = ref.resblocks[i * nk + j]
                xr = r16(x) if i >= x16_from_stage else x
                for l in range(3):
                    a1, a2 = rb.activations[2 * l], rb.activations[2 * l + 1]
                    c1, c2 = rb.convs1[l], rb.convs2[l]
                    xt = r16(a1(xr))
                    xt = F.conv1d(xt, r16(c1.weight), c1.bias, dilation=c1.dilation, padding=c1.padding)
                    if y1_bf16:
                        xt = r16(xt)
                    xt = r16(a2(xt))
                    xt = F.conv1d(xt, r16(c2.weight), c2.bias, padding=c2.padding)
                    xr = xt + xr
                    if i >= res16_from_stage and l < 2:  # inter-layer residual stream of the block stored as bf16
                        xr = r16(xr)
                xs = xr if xs is None else xs + xr
            x = xs / nk
        x = ref.activation_post(x)
        x = ref.conv_post(x)
        return torch.clamp(x, -1, 1)

if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count())
    bigvgan, _, _ = G.import_reference()
    h = bigvgan.load_hparams_from_json("/root/reference/indextts/s2mel/modules/bigvgan/config.json")
    for seed, T in ((1234, 96), (7, 64)):
        torch.manual_seed(seed)
        ref = bigvgan.BigVGAN(h, use_cuda_kernel=False)
        ref.remove_weight_norm(); ref.eval()
        g = torch.Generator().manual_seed(seed)
        with torch.no_grad():
            for p in ref.parameters():
                pass
            # randomise alpha/beta as SURVEY recommends
            for n, p in ref.named_parameters():
                if n.endswith("alpha") or n.endswith("beta"):
                    p.copy_(0.5 * torch.randn(p.shape, generator=g))
        mel = torch.randn(1, 80, T, generator=g)
        with torch.no_grad():
            w32 = ref(mel)
        for y1 in (False, True):
            w = run(ref, mel, y1)
            print(f"seed {seed} T {T} y1_bf16={y1}: SNR {O.snr_db(w32.numpy(), w.numpy()):.2f} dB, max-abs {float((w-w32).abs().max()):.2e}", flush=True)
        # bf16 storage of the blocks' inter-layer residual stream (res16) / of the stage input X (x16) from a given stage on
        for rs, xs_ in ((3, 99), (0, 99), (3, 3), (0, 0)):
            w = run(ref, mel, True, res16_from_stage=rs, x16_from_stage=xs_)
            print(f"seed {seed} T {T} res16>={rs} x16>={xs_}: SNR {O.snr_db(w32.numpy(), w.numpy()):.2f} dB", flush=True)
