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
  cpu_baseline  the CPU oracle port (oracle/torch_port.py: the reference's op sequence on torch CPU kernels) timed on this box's host cores on a bounded sample
--impl reference: the CPU arm only (the reference is Python/torch and cannot travel to the box; its restatement in
oracle/ is timed with all host threads), same metric/config/unit.
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
    """The reference's CPU path restated op for op on torch CPU kernels (oracle/torch_port.py), all host threads,
    on a bounded sample of the workload: one utterance of `frames` mel frames per step."""
    import numpy as np
    import torch
    from oracle import bigvgan_oracle as O, torch_port as TP
    threads = threads or os.cpu_count()
    torch.set_num_threads(threads)
    frames = frames or 431  # 5 s (BASELINE.json configs[0]): 1/32 of the cfg2 batch
    sd = TP.prepare(O.make_state_dict(O.DEFAULT_HPARAMS, seed=1))
    mel = torch.from_numpy(np.random.default_rng(0).standard_normal((1, NUM_MELS, frames)).astype(np.float32))
    for _ in range(warmup):
        TP.forward(sd, mel)
    t0 = time.perf_counter()
    for _ in range(steps):
        TP.forward(sd, mel)
    dt = (time.perf_counter() - t0) / steps
    audio_s = frames * HOP / SR
    return audio_s / dt, dt, threads, (f"1 utterance x {frames} frames ({audio_s:.2f} s audio) per step, fp32, torch CPU "
                                       f"kernels, {threads} threads")


_REAL_STDOUT = None


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    v, dt, threads, sample = cpu_port_audio_s_per_s(steps=max(1, min(args.steps, 5)), warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": "vocoder audio-sec/sec (RTF^-1)", "value": v, "unit": "audio-s/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "CPU arm: bounded sample of the workload"},
        "cpu_baseline": {"value": v, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample},
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

    def step_long():
        with torch.no_grad():
            return sharding.forward_time_sharded(fwd, mel_long, HOP, dst=0)

    for _ in range(2):
        step_long()
    ms_long = timed(step_long, 3)
    long_form = {"workload": "cfg4: one 120 s mel (80x10336) time-sharded over the ranks, 38-frame halo, P2P stitch to rank 0",
                 "value": (long_T * HOP / SR) / (ms_long * 1e-3), "unit": "audio-s/s", "ms": ms_long, "n_gpus": world,
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
    conv = by.get("conv1d", {"ms": 1.0, "flops": 0.0, "n": 0})
    convt = by.get("convtranspose1d", {"ms": 0.0, "flops": 0.0, "n": 0})
    mma_ms = conv["ms"] + convt["ms"]
    mma_flops = conv["flops"] + convt["flops"]
    ach = mma_flops / (mma_ms * 1e-3) / 1e12 if mma_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("conv_umma_kernel_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "tensor", "kernel": "conv_umma_kernel (tcgen05 implicit-GEMM conv + polyphase convT)",
                "achieved": ach, "peak": peaks["tf_sus"], "unit": "TFLOP/s", "frac": ach / peaks["tf_sus"],
                "traffic": traffic, "launches": conv["n"] + convt["n"], "share_of_step": mma_ms / max(sum(d["ms"] for d in by.values()), 1e-9),
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

    cpu = None
    if not args.no_cpu_baseline:
        v, dt, threads, sample = cpu_port_audio_s_per_s(steps=3, warmup=1)
        cpu = {"value": v, "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample}

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
                               "definition": "sum over layers of max(FLOPs/bf16 sustained peak, algorithmic bytes/HBM peak) / sum of CUDA-event kernel times"},
        "kernels": kernels,
        "long_form": long_form,
        "v1_generator": v1_line,
        "latency": latency,
        "cpu_baseline": cpu,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
