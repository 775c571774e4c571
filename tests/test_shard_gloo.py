"""N > 1 host path on CPU: world_size-2 `gloo` process groups (no GPU).

Covers what the multi-GPU run does around the kernels: every rank draws the same global sample sequence, keeps
its disjoint slice (psla_b200/shard.py), builds its own RawBatch with the reference's label logic, and the
bench's aggregation (sum of clips, max-over-ranks time).  There is no collective ON the path; the collectives
here only gather evidence for the assertions.
"""
import json
import os
import socket
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from psla_b200.shard import ShardedBatchSampler, rank_slice, shard_clip_ids
from psla_b200.synth import TINY_CLASSES, TINY_LABELS, tiny_dataset_waves

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_dataset(tmp):
    from psla_b200.dataloaders import AudiosetDataset
    waves = tiny_dataset_waves()
    store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
    jpath, cpath = os.path.join(tmp, "t.json"), os.path.join(tmp, "l.csv")
    if not os.path.exists(jpath):
        json.dump({"data": [{"wav": f"mem://{i}.wav", "labels": TINY_LABELS[i]} for i in range(len(waves))]}, open(jpath, "w"))
        with open(cpath, "w") as f:
            f.write("index,mid,display_name\n" + "".join(f"{i},{m},c{i}\n" for i, m in enumerate(TINY_CLASSES)))
    conf = {"num_mel_bins": 128, "target_length": 256, "freqm": 48, "timem": 192, "mixup": 0.0, "dataset": "audioset",
            "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
    return AudiosetDataset(jpath, conf, cpath, loader=lambda p: (store[p].clone(), 16000), verbose=False), waves


def _global_sequence(n_items, n_draw, seed):
    g = torch.Generator().manual_seed(seed)
    weights = torch.arange(1, n_items + 1, dtype=torch.double)
    return list(torch.utils.data.WeightedRandomSampler(weights, n_draw, replacement=True, generator=g))


def _worker(rank, port, tmp, per_rank_batch, n_draw, out_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(WORLD), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        ds, waves = _make_dataset(tmp)
        seq = _global_sequence(len(ds), n_draw, seed=7)  # same seed on every rank
        sampler = ShardedBatchSampler(seq, per_rank_batch)  # rank / world from the torchrun-style environment
        steps = []
        for batch_ids in sampler:
            batch = ds.collate([ds[i] for i in batch_ids], pinned=False)
            # what this rank would hand to psla_frontend_forward: packed waveforms + per-clip metadata
            assert batch.waves.wav.dtype == torch.float32 and not batch.waves.wav.is_cuda
            assert (batch.waves.offsets % 4 == 0).all()
            for k, i in enumerate(batch_ids):
                n = int(batch.waves.lengths[batch.src1[k]])
                o = int(batch.waves.offsets[batch.src1[k]])
                assert torch.equal(batch.waves.wav[o:o + n], waves[i][0])
            steps.append((batch_ids, batch.labels.sum(1).tolist()))
        # evidence gathering (NOT part of the data path)
        gathered = [None] * WORLD
        dist.all_gather_object(gathered, steps)
        clips = torch.tensor([sum(len(s[0]) for s in steps)], dtype=torch.int64)
        dist.all_reduce(clips, op=dist.ReduceOp.SUM)
        t = torch.tensor([1.0 + rank], dtype=torch.float64)  # bench.py's max-over-ranks timing reduction
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            out_q.put((gathered, int(clips.item()), float(t.item()), seq))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("per_rank_batch,n_draw", [(3, 12), (4, 21)])
def test_two_rank_sharding_covers_the_global_batches_exactly(per_rank_batch, n_draw):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    with tempfile.TemporaryDirectory() as tmp:
        procs = [ctx.Process(target=_worker, args=(r, port, tmp, per_rank_batch, n_draw, q)) for r in range(WORLD)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
            assert p.exitcode == 0
    gathered, clips, tmax, seq = q.get()
    assert clips == n_draw and tmax == float(WORLD)
    g = WORLD * per_rank_batch
    n_steps = max(len(s) for s in gathered)
    rebuilt = []
    for k in range(n_steps):
        for r in range(WORLD):  # rank order inside each global batch
            if k < len(gathered[r]):
                rebuilt += gathered[r][k][0]
    assert rebuilt == [int(i) for i in seq]  # union over ranks == the single-process sequence, in order
    for k in range(n_draw // g):  # full global batches: equal per-rank sizes, disjoint positions
        assert all(len(gathered[r][k][0]) == per_rank_batch for r in range(WORLD))


def test_rank_slice_properties():
    for n in (0, 1, 5, 100, 101, 360, 4096):
        for world in (1, 2, 3, 4, 8):
            sl = [rank_slice(n, r, world) for r in range(world)]
            assert sl[0].start == 0 and sl[-1].stop == n
            assert all(sl[r].stop == sl[r + 1].start for r in range(world - 1))
            sizes = [s.stop - s.start for s in sl]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        rank_slice(10, 2, 2)
    assert shard_clip_ids(list(range(360)), 3, 8) == list(range(135, 180))


def test_sampler_len_and_drop_last():
    seq = list(range(23))
    for world in (1, 2, 4):
        for bs in (1, 3, 5):
            for drop in (False, True):
                per_rank = [list(ShardedBatchSampler(seq, bs, r, world, drop)) for r in range(world)]
                for r in range(world):
                    assert len(per_rank[r]) == len(ShardedBatchSampler(seq, bs, r, world, drop))
                flat = sorted(i for pr in per_rank for b in pr for i in b)
                keep = (len(seq) // (world * bs)) * world * bs if drop else len(seq)
                assert flat == seq[:keep]
