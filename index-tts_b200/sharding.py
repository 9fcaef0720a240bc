"""Multi-GPU partitioning of the vocoder path: one process per GPU, no data-path collective.

The reference is single-GPU and has no distributed code on this path (no NCCL call site); the
two ways the path shards naturally are:

* **utterances** of a batch go to different GPUs (``split_batch`` / ``forward_batch_sharded``) -- weights replicated,
  nothing exchanged during compute;
* **long-form mels** are split along time (``time_shards`` / ``forward_time_sharded``): every rank runs the ordinary
  ``BigVGAN.forward`` on its frame range plus ``RECEPTIVE_HALO_FRAMES`` = 38 frames of *real* neighbour context on
  each interior side (the exact one-sided receptive field of the generator: conv_post 3 + activation_post 5 + six
  stages of (k=11, d=1,3,5) AMP blocks and transposed convs back-propagated to the mel rate), drops the
  ``hop * halo`` samples computed from the halo and the pieces are stitched on the destination rank.  True sequence
  ends keep the reference's own padding (replicate inside Activation1d, zeros in the convs).

The stitch is point-to-point: every rank ``send``s its core waveform piece straight into the slice of the
destination's output buffer (``torch.distributed`` P2P = NCCL send/recv = direct GPU-to-GPU copies over NVLink on a
B200 box).  There is no all-reduce, all-gather or barrier on the data path because none is needed.

``forward_fn`` is any callable ``mel[B, num_mels, T] -> wav[B, 1, T*hop]`` -- the ``BigVGAN`` module in production;
the world_size-2 ``gloo`` tests drive the same code with CPU tensors and the CPU oracle as ``forward_fn``.
"""
from __future__ import annotations

from typing import Callable, List, NamedTuple, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

RECEPTIVE_HALO_FRAMES = 38  # SURVEY.md §8(e); verified bit-exact on the reference's CPU path


class TimeShard(NamedTuple):
    """This rank owns output frames [core_lo, core_hi) and reads mel frames [lo, hi)."""
    core_lo: int
    core_hi: int
    lo: int
    hi: int


def time_shards(n_frames: int, world: int, halo: int = RECEPTIVE_HALO_FRAMES) -> List[TimeShard]:
    """Contiguous frame ranges, sizes differing by at most one; ranks beyond ``n_frames`` get empty shards."""
    if n_frames < 0 or world < 1 or halo < 0:
        raise ValueError(f"time_shards: bad arguments n_frames={n_frames} world={world} halo={halo}")
    out = []
    for r in range(world):
        c0, c1 = (n_frames * r) // world, (n_frames * (r + 1)) // world
        if c1 <= c0:
            out.append(TimeShard(c0, c0, c0, c0))
        else:
            out.append(TimeShard(c0, c1, max(0, c0 - halo), min(n_frames, c1 + halo)))
    return out


def split_batch(n_items: int, world: int) -> List[List[int]]:
    """Utterance b -> rank b mod world (round-robin keeps ragged batches balanced)."""
    return [list(range(r, n_items, world)) for r in range(world)]


def _rank_world(group) -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def _global_rank(group, group_rank: int) -> int:
    return dist.get_global_rank(group, group_rank) if group is not None else group_rank


def run_time_shard(forward_fn: Callable[[torch.Tensor], torch.Tensor], mel: torch.Tensor, shard: TimeShard,
                   hop: int) -> torch.Tensor:
    """Forward over the shard's padded frame range; returns the core piece ``[B, 1, (core_hi-core_lo)*hop]``."""
    if shard.core_hi <= shard.core_lo:
        return mel.new_empty((mel.shape[0], 1, 0), dtype=torch.float32)
    wav = forward_fn(mel[:, :, shard.lo:shard.hi])
    a = (shard.core_lo - shard.lo) * hop
    return wav[:, :, a:a + (shard.core_hi - shard.core_lo) * hop]


def forward_time_sharded(forward_fn: Callable[[torch.Tensor], torch.Tensor], mel: Optional[torch.Tensor], hop: int,
                         n_frames: Optional[int] = None, dst: int = 0, group=None,
                         halo: int = RECEPTIVE_HALO_FRAMES, mel_on_all_ranks: bool = True,
                         num_mels: int = 80, batch: int = 1, device=None,
                         timings: Optional[dict] = None) -> Optional[torch.Tensor]:
    """Vocode one long mel across the ranks of ``group``; the stitched waveform is returned on rank ``dst`` (None
    elsewhere).

    ``mel_on_all_ranks=True``: every rank already holds the full mel (it is tiny -- 3.3 MB for 120 s) and slices its
    own range.  ``False``: only ``dst`` holds it and P2P-sends each rank its padded range first (``n_frames``,
    ``num_mels``, ``batch`` and ``device`` must then be given on the other ranks).

    ``timings`` (measurement only, CUDA tensors): a dict that receives the CUDA events ``t0`` (start), ``t1`` (this
    rank's forward queued) and ``t2`` (stitch done) recorded on the current stream, so that a caller can split the
    step into compute and stitch time per rank (bench.py ``long_form.compute_ms`` / ``stitch_ms``).
    """
    rank, world = _rank_world(group)
    if mel is not None:
        n_frames = mel.shape[-1]
        batch, num_mels, device = mel.shape[0], mel.shape[1], mel.device
    if n_frames is None:
        raise ValueError("forward_time_sharded: n_frames is required on ranks that do not hold the mel")
    shards = time_shards(n_frames, world, halo)
    mine = shards[rank]

    def mark(name):
        if timings is not None and torch.cuda.is_available():
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            timings[name] = ev

    mark("t0")
    if world == 1:
        out1 = run_time_shard(forward_fn, mel, mine, hop)
        mark("t1")
        mark("t2")
        return out1

    # ---- scatter (only when the mel lives on dst alone): P2P sends of the padded ranges --------------------
    if not mel_on_all_ranks:
        if rank == dst:
            reqs = []
            for r, s in enumerate(shards):
                if r != dst and s.hi > s.lo:
                    reqs.append(dist.isend(mel[:, :, s.lo:s.hi].contiguous(), _global_rank(group, r), group=group))
            local = mel[:, :, mine.lo:mine.hi]
            for q in reqs:
                q.wait()
        else:
            local = torch.empty((batch, num_mels, mine.hi - mine.lo), dtype=torch.float32, device=device)
            if mine.hi > mine.lo:
                dist.recv(local, _global_rank(group, dst), group=group)
        piece = run_time_shard(forward_fn, local, TimeShard(mine.core_lo - mine.lo, mine.core_hi - mine.lo, 0,
                                                            mine.hi - mine.lo), hop)
    else:
        piece = run_time_shard(forward_fn, mel, mine, hop)
    mark("t1")

    # ---- stitch: every rank's core piece goes straight into its slice of dst's output ----------------------
    if rank == dst:
        out = torch.empty((batch, 1, n_frames * hop), dtype=torch.float32, device=piece.device)
        reqs, stage = [], []
        for r, s in enumerate(shards):
            if s.core_hi <= s.core_lo:
                continue
            sl = out[:, :, s.core_lo * hop:s.core_hi * hop]
            if r == dst:
                sl.copy_(piece)
            elif batch == 1:
                reqs.append(dist.irecv(sl, _global_rank(group, r), group=group))  # contiguous: lands in place
            else:
                buf = torch.empty((batch, 1, (s.core_hi - s.core_lo) * hop), dtype=torch.float32, device=out.device)
                reqs.append(dist.irecv(buf, _global_rank(group, r), group=group))
                stage.append((sl, buf))
        for q in reqs:
            q.wait()
        for sl, buf in stage:
            sl.copy_(buf)
        mark("t2")
        return out
    if mine.core_hi > mine.core_lo:
        dist.send(piece.contiguous(), _global_rank(group, dst), group=group)
    mark("t2")
    return None


def forward_batch_sharded(forward_fn: Callable, mels: Sequence[torch.Tensor], hop: int,
                          dst: Optional[int] = None, group=None, time_dim: int = -1,
                          indexed: bool = False) -> List[Optional[torch.Tensor]]:
    """Utterance-parallel vocoding of a ragged list of mels ``[num_mels, T_b]`` (every rank holds the list; each mel
    is a few hundred KB).  Rank ``b mod world`` vocodes utterance b.  With ``dst=None`` each rank keeps its own
    results (entries it does not own are None); otherwise the waveforms are P2P-sent to ``dst``.

    v1 generator: the items are channels-last latents ``[T_b, gpt_dim]`` (``time_dim=0``) and every utterance has its own
    speaker embedding, so ``indexed=True`` calls ``forward_fn(b, x)`` with the utterance index."""
    rank, world = _rank_world(group)
    owners = split_batch(len(mels), world)
    out: List[Optional[torch.Tensor]] = [None] * len(mels)
    for b in owners[rank]:
        xb = mels[b].unsqueeze(0)
        out[b] = (forward_fn(b, xb) if indexed else forward_fn(xb))[0]
    if dst is None or world == 1:
        return out
    if rank == dst:
        reqs = []
        for r in range(world):
            if r == dst:
                continue
            for b in owners[r]:
                out[b] = torch.empty((1, mels[b].shape[time_dim] * hop), dtype=torch.float32, device=mels[b].device)
                reqs.append(dist.irecv(out[b], _global_rank(group, r), group=group))
        for q in reqs:
            q.wait()
        return out
    for b in owners[rank]:
        dist.send(out[b].contiguous(), _global_rank(group, dst), group=group)
    return [None] * len(mels)
