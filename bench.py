#!/usr/bin/env python
"""bench.py -- vocoder audio-seconds per wall-second (BASELINE.json metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one BigVGAN forward over one batch of synthetic mels.  Workload at every N: BASELINE.json configs[1],
batch 16 x 10 s mels (80 x 862 frames) per GPU, bf16 tensor-core mode, random-init weights of
bigvgan_v2_22khz_80band_256x (weak scaling: utterances of a batch go to different GPUs, no data-path collective).

Printed JSON line (rank 0):
  value      audio-s/s with the mels already resident in HBM (CUDA events around each step, max over ranks)
  e2e        same metric through the public module call with HOST mels: pinned H2D copy + forward + D2H of the wav
  roofline   the dominant kernel (tcgen05 implicit-GEMM conv): algorithmic FLOPs of its launches / their CUDA-event
             time, against the measured bf16 peak in MEASURED_PEAKS.json
  cpu_baseline  the UNMODIFIED reference BigVGAN.forward (baseline/_ref, installed by baseline/install_ref.sh; torch CPU
             kernels, all host threads) on a bounded sample: two of the batch's 16 utterances.  Falls back to the oracle
             port (oracle/torch_port.py, kind "port") when baseline/_ref is absent
  parity     SNR / max-abs of the TIMED output tensor against those two reference runs (same weights)
  fp32_mode  the fp32 parity mode (bf16x3 split operands on tcgen05): cfg1-shape latency, its roofline fractions and
             the waveform error of utterance 0 against the same reference run
--impl reference: the CPU arm only, K timed steps after W warm-up steps, one 10 s utterance of the batch per step
(utterances are independent, so audio-s/s of the sample is the metric of the workload); same metric/config/unit.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SR, HOP, NUM_MELS = 22050, 256, 80
CFG2_B, CFG2_T = 16, 862
WORKLOAD = "cfg2: bigvgan_v2_22khz_80band_256x random-init, batch 16 x 10 s synthetic mels (80x862) per GPU"


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), tf=float(d["bf16_tflops"]), tf_sus=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf=1590.0, tf_sus=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        # samples under load = upper half of the distribution
        sm.sort()
        under = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": statistics.median(under) if under else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def make_state_dict_torch(seed=1234):
    """Random-init weights of the named architecture: the module's own seeded construction (identical to the
    reference's, tests/test_abi.py), alpha/beta randomised N(0, 0.5) so per-channel indexing is exercised."""
    import torch
    from index_tts_b200 import bigvgan as B
    h = B.load_hparams_from_json(os.path.join(ROOT, "index-tts_b200", "config.json"))
    torch.manual_seed(seed)
    m = B.BigVGAN(h, use_cuda_kernel=True)
    g = torch.Generator().manual_seed(4321)
    with torch.no_grad():
        for n, p in m.named_parameters():
            if n.endswith(".alpha") or n.endswith(".beta"):
                p.copy_(0.5 * torch.randn(p.shape, generator=g))
    m.remove_weight_norm()
    return m.eval()


def cpu_port_audio_s_per_s(steps=1, warmup=1, frames=None, threads=None):
    """Fallback CPU arm: the reference's CPU path restated op for op on torch CPU kernels (oracle/torch_port.py), all host
    threads, on a bounded sample of the workload: one utterance of `frames` mel frames per step."""
    import numpy as np
    import torch
    from oracle import bigvgan_oracle as O, torch_port as TP
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    frames = frames or CFG2_T
    sd = TP.prepare(O.make_state_dict(O.DEFAULT_HPARAMS, seed=1))
    mel = torch.from_numpy(np.random.default_rng(0).standard_normal((1, NUM_MELS, frames)).astype(np.float32))
    for _ in range(warmup):
        TP.forward(sd, mel)
    t0 = time.perf_counter()
    for _ in range(steps):
        TP.forward(sd, mel)
    dt = (time.perf_counter() - t0) / steps
    audio_s = frames * HOP / SR
    return audio_s / dt, dt, threads, (f"1 of the batch's 16 utterances ({frames} frames, {audio_s:.2f} s audio) per step, fp32, "
                                       f"oracle/torch_port.py on torch CPU kernels, {threads} threads"), "port"


def reference_cpu_model(state_dict=None):
    """The UNMODIFIED reference module (baseline/_ref, installed by baseline/install_ref.sh) on the CPU with the bench
    weights; None when baseline/_ref is absent."""
    import contextlib
    import torch
    from baseline import ref_loader
    if not ref_loader.available():
        return None
    with contextlib.redirect_stdout(sys.stderr):
        bigvgan, _, _ = ref_loader.import_reference()
        h = bigvgan.load_hparams_from_json(os.path.join(ROOT, "index-tts_b200", "config.json"))
        torch.manual_seed(1234)
        m = bigvgan.BigVGAN(h, use_cuda_kernel=False)
        m.remove_weight_norm()
        if state_dict is not None:
            m.load_state_dict(state_dict, strict=True)
    return m.eval()


def cpu_reference_audio_s_per_s(steps=1, warmup=1, frames=None, threads=None, model=None, mels=None):
    """CPU arm: the reference's own BigVGAN.forward (bigvgan.py:360-386, torch path, fp32) on the box's host cores, on a
    bounded sample of the workload -- ONE of the batch's 16 independent utterances per step (same length, same weights).
    Returns (audio-s/s, s/step, threads, sample text, kind, outputs)."""
    import torch
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    frames = frames or CFG2_T
    if model is None:
        model = reference_cpu_model()
    if model is None:
        return cpu_port_audio_s_per_s(steps, warmup, frames, threads) + (None,)
    if mels is None:
        g = torch.Generator().manual_seed(1)
        mels = [torch.randn(1, NUM_MELS, frames, generator=g) for _ in range(max(steps, 1))]
    outs = []
    with torch.inference_mode():
        for i in range(warmup):
            model(mels[i % len(mels)])
        t0 = time.perf_counter()
        for i in range(steps):
            outs.append(model(mels[i % len(mels)]))
        dt = (time.perf_counter() - t0) / steps
    audio_s = frames * HOP / SR
    sample = (f"1 of the batch's 16 utterances ({frames} frames, {audio_s:.2f} s audio) per step, fp32, the unmodified "
              f"reference BigVGAN.forward (baseline/_ref) on torch CPU kernels, {threads} threads")
    return audio_s / dt, dt, threads, sample, "reference", outs


_REAL_STDOUT = None


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # every step really runs: K timed steps after W warm-up steps.  A step is one utterance of the cfg2 batch (862 frames,
    # 10 s); on a slow host the utterance is shortened (431 / 216 / 108 frames -- the sample text says which) so that the
    # whole K + W run stays within ~3 minutes, calibrated by one untimed 108-frame forward.
    W = max(0, args.warmup)
    K = max(1, args.steps)
    import torch
    torch.set_num_threads(os.cpu_count())
    model = reference_cpu_model()
    frames = CFG2_T
    if model is not None:
        with torch.inference_mode():
            probe = torch.randn(1, NUM_MELS, 108)
            model(probe)
            t0 = time.perf_counter()
            model(probe)
            per_frame = (time.perf_counter() - t0) / 108
        budget_s = float(os.environ.get("BVG_REF_BUDGET_S", "180"))
        while frames > 108 and per_frame * frames * (K + W) > budget_s:
            frames = (frames + 1) // 2
    v, dt, threads, sample, kind, _ = cpu_reference_audio_s_per_s(steps=K, warmup=W, frames=frames, model=model)
    line = {
        "impl": "reference", "metric": "vocoder audio-sec/sec (RTF^-1)", "value": v, "unit": "audio-s/s", "n_gpus": args.gpus,
        "steps": max(1, args.steps), "warmup": W, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU arm: each step is a bounded sample of the workload (one of the 16 "
                   "independent utterances of the batch); audio-s/s is per-utterance work / time, so it is the same metric",
                   "runs_on": "rank 0 only, host cores; the other ranks exit"},
        "cpu_baseline": {"value": v, "unit": "audio-s/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=CFG2_B)
    ap.add_argument("--frames", type=int, default=CFG2_T)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-v1", action="store_true", help="skip the v1-generator leg")
    ap.add_argument("--layer-table", default=None, help="write the per-launch table (JSON) to this path")
    args = ap.parse_args()
    # stdout carries exactly ONE JSON line: everything else that writes to fd 1 (NCCL's version banner, the reference's
    # prints) is sent to stderr, and the line itself goes to the saved descriptor at the end
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from index_tts_b200 import bigvgan as B, ops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: libbvg has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = max(3, args.warmup)

    import contextlib
    with contextlib.redirect_stdout(sys.stderr):  # the reference's remove_weight_norm() prints; stdout carries ONE JSON line
        model = make_state_dict_torch().to(dev)
    model.set_precision(args.precision)
    Bsz, T = args.batch, args.frames
    gen = torch.Generator().manual_seed(1 + rank)
    mel_host = torch.randn(Bsz, NUM_MELS, T, generator=gen).pin_memory()
    mel_dev = mel_host.to(dev)
    wav_host = torch.empty(Bsz, 1, T * HOP).pin_memory()
    audio_s = Bsz * T * HOP / SR
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier()
        for e0, e1 in evs:
            flush.fill_(1)            # L2 flush between timed iterations (untimed)
            e0.record()
            fn()
            e1.record()
        barrier()
        ms = sum(e0.elapsed_time(e1) for e0, e1 in evs)
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    def step_resident():
        with torch.no_grad():
            return model(mel_dev)

    def step_e2e():
        with torch.no_grad():
            m = mel_host.to(dev, non_blocking=True)
            w = model(m)
            wav_host.copy_(w, non_blocking=True)

    for _ in range(W):
        step_resident()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ops.reset_launch_count()
    ms_step = timed(step_resident, args.steps)
    launches = ops.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps)

    # ---- long-form leg (BASELINE.json configs[3]): one 120 s mel split along time over the N ranks with the
    # 38-frame receptive-field halo, core pieces P2P-sent (NCCL send/recv) into rank 0's output buffer.  No collective
    # on the data path; the all_reduce(MAX) below only reduces the timing.
    from index_tts_b200 import sharding
    long_T = 10336
    mel_long = torch.randn(1, NUM_MELS, long_T, generator=torch.Generator().manual_seed(2)).to(dev)
    fwd = lambda m: model(m)

    long_tm = {}

    def step_long():
        with torch.no_grad():
            return sharding.forward_time_sharded(fwd, mel_long, HOP, dst=0, timings=long_tm)

    for _ in range(2):
        step_long()
    ms_long = timed(step_long, 3)
    # compute / stitch split of one more (untimed-as-a-whole) step, per rank, max over ranks
    barrier()
    wav_long = step_long()
    torch.cuda.synchronize()
    split = torch.tensor([long_tm["t0"].elapsed_time(long_tm["t1"]), long_tm["t1"].elapsed_time(long_tm["t2"])], device=dev)
    if world > 1:
        dist.all_reduce(split, op=dist.ReduceOp.MAX)
    # parity of the stitch: rank 0 also runs the UNSHARDED forward of the same mel once (untimed) and compares
    long_err = None
    if rank == 0:
        with torch.no_grad():
            ref_long = model(mel_long)
        long_err = float((wav_long - ref_long).abs().max().item())
        del ref_long
    long_form = {"workload": "cfg4: one 120 s mel (80x10336) time-sharded over the ranks, 38-frame halo, P2P stitch to rank 0",
                 "value": (long_T * HOP / SR) / (ms_long * 1e-3), "unit": "audio-s/s", "ms": ms_long, "n_gpus": world,
                 "compute_ms": float(split[0].item()), "stitch_ms": float(split[1].item()),
                 "max_abs_vs_unsharded": long_err,
                 "redundant_frames_per_rank": 0 if world == 1 else 2 * sharding.RECEPTIVE_HALO_FRAMES,
                 "exchange": "none" if world == 1 else "NCCL send/recv of core waveform pieces (no collective)"}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = load_peaks()
    # per-launch table of the same plan, live, CUDA events on the launching stream
    for _ in range(2):
        B.profile_forward(model, mel_dev)
    _, recs = B.profile_forward(model, mel_dev)
    by = {}
    roof_ms_total, meas_ms_total = 0.0, 0.0
    for r in recs:
        d = by.setdefault(r["kind"], {"ms": 0.0, "flops": 0.0, "bytes": 0.0, "n": 0, "roof_ms": 0.0})
        roof = max(r["flops"] / (peaks["tf_sus"] * 1e12), r["bytes"] / (peaks["hbm"] * 1e9)) * 1e3
        d["ms"] += r["ms"]; d["flops"] += r["flops"]; d["bytes"] += r["bytes"]; d["n"] += 1; d["roof_ms"] += roof
        if r["kind"] != "layout":
            roof_ms_total += roof
            meas_ms_total += r["ms"]
    # The same roofline with the bytes the plan's kernels really move in bf16 mode (fp32 residual stream):
    # SURVEY.md 8(d) charges 2 bytes per element everywhere, but the second conv of every AMP sub-block reads an fp32
    # residual and writes an fp32 result (bf16 operand 2 + 4 + 4 bytes per output element), and the first activation of a
    # sub-block reads the fp32 residual stream (4 + 2 bytes).  Layers are recognised from the record sequence.
    roof_actual_ms = 0.0
    if args.precision == "bf16":
        for i, r in enumerate(recs):
            if r["kind"] == "layout":
                continue
            rows, cin, cout, k = float(r["rows"]), float(r["cin"]), float(r["cout"]), float(max(r["k"], 1))
            b_act = r["bytes"]
            if r["kind"] == "conv1d" and r["stage"] >= 0:
                plain = 2.0 * rows * (cin + cout) + 2.0 * cin * cout * k + 4.0 * cout
                if r["bytes"] > plain * 1.05:  # carries the residual term: fp32 residual in, fp32 out
                    b_act = 2.0 * rows * cin + 8.0 * rows * cout + 2.0 * cin * cout * k + 4.0 * cout
            elif r["kind"] == "activation1d":
                heads = max(1, int(round(r["flops"] / (62.0 * rows * cin))))
                nxt = recs[i + 1] if i + 1 < len(recs) else None
                if r["stage"] < 0:
                    b_act = 8.0 * rows * cin                       # activation_post: fp32 in, fp32 out
                elif heads > 1:
                    b_act = (4.0 + 2.0 * heads) * rows * cin       # shared up-sampler: one fp32 read, `heads` bf16 outputs
                elif nxt is not None and nxt["kind"] == "conv1d" and nxt["bytes"] <= (2.0 * rows * 2 * cin + 2.0 * cin * cin * max(nxt["k"], 1) + 4.0 * cin) * 1.05:
                    b_act = 6.0 * rows * cin                       # first activation of a sub-block: fp32 residual stream in
            elif r["kind"] == "convtranspose1d":
                b_act = r["bytes"] + 2.0 * rows * cout             # the up-sampled residual stream is written as fp32 (4 B, not 2)
            roof_actual_ms += max(r["flops"] / (peaks["tf_sus"] * 1e12), b_act / (peaks["hbm"] * 1e9)) * 1e3
    conv = by.get("conv1d", {"ms": 1.0, "flops": 0.0, "n": 0})
    convt = by.get("convtranspose1d", {"ms": 0.0, "flops": 0.0, "n": 0})
    mma_ms = conv["ms"] + convt["ms"]
    mma_flops = conv["flops"] + convt["flops"]
    ach = mma_flops / (mma_ms * 1e-3) / 1e12 if mma_ms > 0 else 0.0
    # DRAM traffic of the dominant kernel from ncu (profiles/ncu_traffic.json, written by tools/ncu_summary.py):
    # `traffic` is the per-launch AVERAGE of dram__bytes_read.sum + dram__bytes_write.sum over every conv_umma_kernel
    # launch of one step (same aggregation as `achieved`), when a launch-list capture with those metrics exists;
    # `traffic_one_launch` is the figure of the single --set full capture (one stage-1 k = 11 layer).
    traffic, traffic_one = None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        try:
            tj = json.load(open(tpath))
            traffic = tj.get("conv_umma_kernel_bytes_per_launch_avg")
            traffic_one = {"bytes": tj.get("conv_umma_kernel_bytes_per_launch"), "layer": tj.get("layer"),
                           "algorithmic_bytes": tj.get("algorithmic_bytes")}
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "conv_umma_kernel (tcgen05 implicit-GEMM conv + polyphase convT)",
                "achieved": ach, "peak": peaks["tf_sus"], "unit": "TFLOP/s", "frac": ach / peaks["tf_sus"],
                "traffic": traffic, "traffic_one_launch": traffic_one,
                "algorithmic_bytes_per_launch_avg": (conv["bytes"] + convt["bytes"]) / max(conv["n"] + convt["n"], 1),
                "launches": conv["n"] + convt["n"], "share_of_step": mma_ms / max(sum(d["ms"] for d in by.values()), 1e-9),
                "peak_source": peaks["source"] + ", sustained bf16 (kernel timed inside a long step)"}
    act = by.get("activation1d")
    kernels = {k: {"ms": round(d["ms"], 4), "launches": d["n"], "roofline_ms": round(d["roof_ms"], 4),
                   "frac_of_per_layer_roofline": (d["roof_ms"] / d["ms"]) if d["ms"] > 0 else None} for k, d in by.items()}
    if act:
        kernels["activation1d"]["achieved_GBs_algorithmic"] = act["bytes"] / (act["ms"] * 1e-3) / 1e9
        kernels["activation1d"]["hbm_peak_GBs"] = peaks["hbm"]
    if args.layer_table:
        os.makedirs(os.path.dirname(os.path.abspath(args.layer_table)), exist_ok=True)
        json.dump({"workload": WORKLOAD, "batch": Bsz, "frames": T, "peaks": peaks, "records": recs}, open(args.layer_table, "w"), indent=1)

    # ---- v1 generator leg (reference indextts/BigVGAN/models.py): the conditioned 24 kHz / 1024x
    # generator through its public forward(x, mel_ref) -- ECAPA-TDNN speaker encoder (torch) + bvg_forward_cond -- on
    # 16 x 10 s of synthetic GPT latents, and the decoder alone with the embedding given.
    v1_line = None
    if not args.no_v1 and world == 1:  # timed() barriers across ranks; the other ranks have left by now
        from index_tts_b200 import v1 as V1
        hv = B.AttrDict(dict(V1.DEFAULT_HPARAMS))
        torch.manual_seed(4321)
        mv = V1.BigVGAN(hv, use_cuda_kernel=True)
        with torch.no_grad():
            for n_, p_ in mv.named_parameters():
                if n_.endswith(".alpha") or n_.endswith(".beta"):
                    p_.normal_(0.0, 0.5)
        mv.remove_weight_norm()
        mv = mv.to(dev).eval().set_precision(args.precision)
        Tv = 235  # 235 latent frames x 1024 / 24 kHz = 10.03 s
        gv = torch.Generator().manual_seed(3)
        xv = torch.randn(Bsz, Tv, hv.gpt_dim, generator=gv).to(dev)
        melv = torch.randn(Bsz, 300, hv.num_mels, generator=gv).to(dev)
        with torch.no_grad():
            embv = mv.speaker_encoder(melv)
        fv = lambda: mv(xv, melv)
        fd = lambda: mv.forward_with_embedding(xv, embv)
        with torch.no_grad():
            for _ in range(3):
                fv()
            ms_v1 = timed(fv, 3)
            fd()
            ms_v1d = timed(fd, 3)
        for _ in range(2):
            mv.profile_with_embedding(xv, embv)
        _, recs_v1 = mv.profile_with_embedding(xv, embv)
        peaks_v = load_peaks()
        roof_v1 = sum(max(r["flops"] / (peaks_v["tf_sus"] * 1e12), r["bytes"] / (peaks_v["hbm"] * 1e9)) * 1e3
                      for r in recs_v1 if r["kind"] != "layout")
        meas_v1 = sum(r["ms"] for r in recs_v1 if r["kind"] != "layout")
        audio_v1 = Bsz * Tv * 1024 / 24000.0
        v1_line = {"workload": "v1 generator (models.py shape class: 1536 ch, rates 4,4,4,4,2,2, gpt_dim 1024, D 512), "
                               f"batch {Bsz} x {Tv} latent frames (10 s each), random-init, {args.precision}",
                   "value": audio_v1 / (ms_v1 * 1e-3), "unit": "audio-s/s", "ms_per_step": ms_v1,
                   "decoder_only_ms": ms_v1d, "decoder_only_value": audio_v1 / (ms_v1d * 1e-3),
                   "per_layer_roofline": {"roofline_ms": roof_v1, "measured_ms": meas_v1,
                                          "frac": roof_v1 / meas_v1 if meas_v1 > 0 else None, "launches": len(recs_v1)}}
        del mv, xv, melv

    # ---- single-utterance latency (BASELINE.json configs[0] shape: one 5 s mel, 80x431; the reference pipeline calls the
    # vocoder with B = 1, infer_v2.py:860): eager launch plan vs CUDA-graph replay (BigVGAN.set_cuda_graphs)
    latency = None
    if world == 1:
        mel1 = torch.randn(1, NUM_MELS, 431, generator=torch.Generator().manual_seed(0)).to(dev)
        f1 = lambda: model(mel1)
        with torch.no_grad():
            for _ in range(3):
                f1()
            ms_eager = timed(f1, 10)
            model.set_cuda_graphs(4096)
            for _ in range(3):
                f1()
            ms_graph = timed(f1, 10)
            model.set_cuda_graphs(0)
        latency = {"workload": "cfg1 shape: 1 x 5 s mel (80x431), " + args.precision, "eager_ms": ms_eager,
                   "cuda_graph_ms": ms_graph, "audio_s_per_s_graph": (431 * HOP / SR) / (ms_graph * 1e-3)}

    # ---- cfg3 point (BASELINE.json configs[2]): the public [B, C, T] Activation1d op (north_star (a)) at C = 1536,
    # T = 2^21, both I/O dtypes, isolated, L2 flushed between launches; algorithmic bytes = one read + one write.
    act_op = None
    if world == 1:
        act_op = {"workload": "cfg3: Activation1d [1, 1536, 2^21], isolated launch, 256 MiB L2 flush between launches",
                  "hbm_peak_GBs": peaks["hbm"], "points": []}
        Cc, Tc = 1536, 1 << 21
        alc = 0.5 * torch.randn(Cc, device=dev)
        bec = 0.5 * torch.randn(Cc, device=dev)
        # sine modes: "fast" = MUFU cos (what the bf16 forward uses), "precise" = range-reduced sinf (fp32 parity mode)
        for dt, name, prec_c in ((torch.float32, "f32", False), (torch.float32, "f32", True), (torch.bfloat16, "bf16", False)):
            xc = torch.randn(1, Cc, Tc, device=dev).to(dt)
            fc = lambda: ops.activation1d(xc, alc, bec, prec_c)
            for _ in range(2):
                yc = fc()
            ms_c = timed(fc, 5)
            nbytes = 2.0 * xc.numel() * xc.element_size()
            act_op["points"].append({"dtype": name, "sine": "precise" if prec_c else "fast", "ms": ms_c, "algorithmic_GBs": nbytes / (ms_c * 1e-3) / 1e9,
                                     "frac_of_hbm_peak": nbytes / (ms_c * 1e-3) / 1e9 / peaks["hbm"],
                                     "Gelem_per_s": xc.numel() / (ms_c * 1e-3) / 1e9})
            del xc, yc
        torch.cuda.empty_cache()

    # ---- CPU baseline + parity of the TIMED tensor: the reference's own forward (fp32, host cores, same weights) on two
    # of the 16 utterances of the timed batch; the same two runs are the timed CPU baseline (bounded sample).
    cpu, parity, outs = None, None, None
    if not args.no_cpu_baseline:
        import numpy as np
        with contextlib.redirect_stdout(sys.stderr):
            sd_cpu = {k: v.detach().float().cpu() for k, v in model.state_dict().items()}
            ref_m = reference_cpu_model(sd_cpu)
        picks = [0, Bsz - 1] if Bsz > 1 else [0]
        mels_cpu = [mel_host[i:i + 1].clone() for i in picks]
        v, dt, threads, sample, kind, outs = cpu_reference_audio_s_per_s(steps=len(picks), warmup=1, frames=T, model=ref_m,
                                                                         mels=mels_cpu)
        cpu = {"value": v, "unit": "audio-s/s", "cores": threads, "kind": kind, "sample": sample}
        if outs is not None:
            with torch.no_grad():
                wav_timed = step_resident()  # the tensor the timed loop produces (same inputs, same plan)
            snrs, errs = [], []
            for i, o in zip(picks, outs):
                ours = wav_timed[i].float().cpu().numpy().astype(np.float64).ravel()
                refw = o[0].numpy().astype(np.float64).ravel()
                err = ours - refw
                snrs.append(10.0 * float(np.log10((refw ** 2).sum() / max((err ** 2).sum(), 1e-300))))
                errs.append(float(np.abs(err).max()))
            bar = 40.0 if args.precision == "bf16" else 80.0
            parity = {"against": "the unmodified reference BigVGAN.forward (baseline/_ref), fp32 on CPU, same weights",
                      "utterances": picks, "snr_db": min(snrs), "snr_db_each": snrs, "max_abs": max(errs),
                      "bar_snr_db": bar, "ok": bool(min(snrs) >= bar and (args.precision == "bf16" or max(errs) <= 1e-4))}

    # ---- fp32 parity mode (the precision the reference runs the path in, infer_v2.py:830-831, 860): bf16x3 split
    # operands on the tcgen05 conv kernel (BVG_F32_SPLIT_TC).  cfg1 shape latency, its per-layer roofline (a conv's
    # tensor-core ceiling is the bf16 peak / 6 partial products; bytes at 4 B per element) and the waveform error
    # against the unmodified reference on utterance 0 of the timed batch.
    fp32_mode = None
    if world == 1:
        model.set_precision("fp32")
        mel1 = torch.randn(1, NUM_MELS, 431, generator=torch.Generator().manual_seed(0)).to(dev)
        f1 = lambda: model(mel1)
        with torch.no_grad():
            for _ in range(3):
                f1()
            ms32_eager = timed(f1, 10)
            model.set_cuda_graphs(4096)
            for _ in range(3):
                f1()
            ms32_graph = timed(f1, 10)
            model.set_cuda_graphs(0)
        for _ in range(2):
            B.profile_forward(model, mel1)
        _, recs32 = B.profile_forward(model, mel1)
        tf32_peak = peaks["tf_sus"] / 6.0
        roof32 = sum(max(r["flops"] / (tf32_peak * 1e12), r["bytes"] / (peaks["hbm"] * 1e9)) * 1e3
                     for r in recs32 if r["kind"] != "layout")
        meas32 = sum(r["ms"] for r in recs32 if r["kind"] != "layout")
        conv32 = [r for r in recs32 if r["kind"] in ("conv1d", "convtranspose1d")]
        conv32_ms = sum(r["ms"] for r in conv32)
        ach32 = sum(r["flops"] for r in conv32) / (conv32_ms * 1e-3) / 1e12 if conv32_ms > 0 else 0.0
        par32 = None
        if cpu is not None and parity is not None:
            import numpy as np
            with torch.no_grad():
                w32 = model(mel_dev[0:1])[0].float().cpu().numpy().astype(np.float64).ravel()
            refw = outs[0][0].numpy().astype(np.float64).ravel()
            e32 = w32 - refw
            par32 = {"utterance": 0, "max_abs": float(np.abs(e32).max()),
                     "snr_db": 10.0 * float(np.log10((refw ** 2).sum() / max((e32 ** 2).sum(), 1e-300)))}
            par32["ok"] = bool(par32["max_abs"] <= 1e-4 and par32["snr_db"] >= 80.0)
        fp32_mode = {"workload": "cfg1 shape: 1 x 5 s mel (80x431), fp32 parity mode (bf16x3 split operands on tcgen05)",
                     "eager_ms": ms32_eager, "cuda_graph_ms": ms32_graph,
                     "audio_s_per_s_graph": (431 * HOP / SR) / (ms32_graph * 1e-3),
                     "roofline": {"bound": "tensor", "achieved": ach32, "peak": tf32_peak, "unit": "TFLOP/s (fp32-equivalent)",
                                  "frac": ach32 / tf32_peak, "peak_source": "sustained bf16 peak / 6 partial products"},
                     "per_layer_roofline": {"roofline_ms": roof32, "measured_ms": meas32,
                                            "frac": roof32 / meas32 if meas32 > 0 else None, "launches": len(recs32)},
                     "parity": par32}
        model.set_precision(args.precision)

    line = {
        "metric": "vocoder audio-sec/sec (RTF^-1)", "value": world * audio_s / (ms_step * 1e-3), "unit": "audio-s/s",
        "n_gpus": world, "steps": args.steps, "warmup": W, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch_per_gpu": Bsz, "mel_frames": T, "audio_s_per_step_per_gpu": audio_s,
                   "precision": args.precision, "l2": "256 MiB flush write between timed iterations (untimed)",
                   "sharding": "batch items across GPUs, no collective"},
        "e2e": {"value": world * audio_s / (ms_e2e * 1e-3), "unit": "audio-s/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": mel_host.numel() * 4, "d2h_bytes_per_step": wav_host.numel() * 4},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "per_layer_roofline": {"roofline_ms": roof_ms_total, "measured_ms": meas_ms_total,
                               "frac": roof_ms_total / meas_ms_total if meas_ms_total > 0 else None,
                               "definition": "sum over layers of max(FLOPs/bf16 sustained peak, algorithmic bytes/HBM peak) / sum of CUDA-event kernel times",
                               "roofline_ms_actual_traffic": roof_actual_ms if roof_actual_ms > 0 else None,
                               "frac_actual_traffic": (roof_actual_ms / meas_ms_total) if (roof_actual_ms > 0 and meas_ms_total > 0) else None,
                               "actual_traffic_note": "same sum with the bytes the bf16-mode plan really moves (fp32 residual stream: 10 B per element for the second conv of a sub-block, 6 B for its first activation) instead of SURVEY.md 8(d)'s 2 B per element"},
        "kernels": kernels,
        "long_form": long_form,
        "v1_generator": v1_line,
        "latency": latency,
        "fp32_mode": fp32_mode,
        "activation1d_op": act_op,
        "cpu_baseline": cpu,
        "parity": parity,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
