"""Multi-GPU host logic (index-tts_b200/sharding.py) on CPU: plans, and the world_size-2 gloo path with the CPU oracle
standing in for the vocoder (the GPU box runs the same code over NCCL with the real module)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import MINI_H, O

from index_tts_b200 import sharding as S

HOP = 256


def test_time_shard_plan_covers_every_frame_once():
    for T, W in ((10336, 8), (10336, 2), (431, 4), (100, 3), (5, 8), (0, 2), (77, 1)):
        sh = S.time_shards(T, W)
        assert len(sh) == W
        covered = []
        for s in sh:
            assert s.lo <= s.core_lo <= s.core_hi <= s.hi <= max(T, 0)
            if s.core_hi > s.core_lo:
                assert s.core_lo - s.lo == min(S.RECEPTIVE_HALO_FRAMES, s.core_lo)          # real context, never padding
                assert s.hi - s.core_hi == min(S.RECEPTIVE_HALO_FRAMES, T - s.core_hi)
            covered += list(range(s.core_lo, s.core_hi))
        assert covered == list(range(T))
    assert S.time_shards(10336, 8)[3] == (3876, 5168, 3838, 5206)
    with pytest.raises(ValueError):
        S.time_shards(10, 0)


def test_batch_split_round_robin():
    assert S.split_batch(16, 8) == [[r, r + 8] for r in range(8)]
    assert S.split_batch(3, 4) == [[0], [1], [2], []]
    assert sorted(sum(S.split_batch(13, 4), [])) == list(range(13))


def _oracle_fn(sd):
    def fn(mel):
        return torch.from_numpy(O.forward(sd, mel.numpy(), MINI_H, np.float32))
    return fn


def test_single_process_time_sharding_equals_unsharded():
    sd = O.make_state_dict(MINI_H, seed=3)
    mel = torch.from_numpy(np.random.default_rng(5).standard_normal((1, 80, 90)).astype(np.float32))
    fn = _oracle_fn(sd)
    full = fn(mel)
    pieces = [S.run_time_shard(fn, mel, s, HOP) for s in S.time_shards(90, 3)]
    assert torch.allclose(torch.cat(pieces, -1), full, atol=1e-6)
    assert torch.equal(S.forward_time_sharded(fn, mel, HOP), full)  # world 1: the ordinary forward


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, scatter, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        sd = O.make_state_dict(MINI_H, seed=3)
        fn = _oracle_fn(sd)
        rng = np.random.default_rng(5)
        mel = torch.from_numpy(rng.standard_normal((1, 80, 90)).astype(np.float32))
        if scatter:  # only rank 1 holds the mel; the result must land on rank 1
            out = S.forward_time_sharded(fn, mel if rank == 1 else None, HOP, n_frames=90, dst=1,
                                         mel_on_all_ranks=False, num_mels=80, batch=1, device="cpu")
            owner = 1
        else:
            out = S.forward_time_sharded(fn, mel, HOP, dst=0)
            owner = 0
        mels = [torch.from_numpy(rng.standard_normal((80, t)).astype(np.float32)) for t in (7, 12, 9)]
        wavs = S.forward_batch_sharded(fn, mels, HOP, dst=0)
        # v1 generator: channels-last latents, one speaker embedding per utterance (forward_fn gets the index)
        hv = dict(O.V1_HPARAMS, upsample_initial_channel=64, gpt_dim=16, speaker_embedding_dim=8, upsample_rates=[4, 2],
                  upsample_kernel_sizes=[8, 4], resblock_kernel_sizes=[3], resblock_dilation_sizes=[[1, 3]])
        sdv = O.make_state_dict(hv, seed=4)
        lat = [torch.from_numpy(rng.standard_normal((t, 16)).astype(np.float32)) for t in (5, 8, 6)]
        emb = rng.standard_normal((3, 8)).astype(np.float32)
        fnv = lambda b, x: torch.from_numpy(O.forward_v1(sdv, x.numpy(), emb[b:b + 1], hv, np.float32))
        wv = S.forward_batch_sharded(fnv, lat, 8, dst=0, time_dim=0, indexed=True)
        if rank == 0:
            errv = max(float((w - fnv(b, x.unsqueeze(0))[0]).abs().max()) for b, (w, x) in enumerate(zip(wv, lat)))
            assert errv <= 1e-6 and tuple(w.shape[-1] for w in wv) == (40, 64, 48), (errv, [w.shape for w in wv])
        else:
            assert wv == [None] * 3
        if rank == owner:
            q.put(("time", float((out - fn(mel)).abs().max()), tuple(out.shape)))
        else:
            assert out is None
        if rank == 0:
            err = max(float((w - fn(m.unsqueeze(0))[0]).abs().max()) for w, m in zip(wavs, mels))
            q.put(("batch", err, tuple(w.shape[-1] for w in wavs)))
        else:
            assert wavs == [None] * 3
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("scatter", [False, True])
def test_world2_gloo_time_and_batch_sharding(scatter):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, scatter, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = {}
    for _ in range(2):
        kind, err, shape = q.get(timeout=240)
        got[kind] = (err, shape)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert got["time"][1] == (1, 1, 90 * HOP) and got["time"][0] <= 1e-6
    assert got["batch"][1] == (7 * HOP, 12 * HOP, 9 * HOP) and got["batch"][0] <= 1e-6
