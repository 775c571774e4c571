"""Data-parallel sharding of the frontend: one process per GPU, disjoint clips per rank, NO collective on the path.

The reference scales with `nn.DataParallel` (src/traintest.py:44-45): ONE process computes every clip's
features on CPU workers, then scatters the [B, T, 128] batch over the GPUs.  Clips are independent units
(SURVEY.md 8(e)), so here every rank draws the SAME global index sequence (same sampler seed, exactly the
sequence the reference's single `DataLoader` would see, `src/run.py:82-94`) and keeps a contiguous slice
`[r*B, (r+1)*B)` of every global batch of `world*B` items.  Mixup partners are arbitrary dataset items loaded
by the same rank (audioset_dataset.py:153), so there is no cross-rank data dependency either.

Host logic only (no CUDA): tested with world_size-2 `gloo` process groups in tests/test_shard_gloo.py.
"""
from __future__ import annotations

import os
from typing import Iterable, Iterator, List, Optional, Sequence, Tuple


def dist_env() -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment; (0, 1, 0) when not launched by it."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def rank_slice(n_items: int, rank: int, world: int) -> slice:
    """Contiguous, balanced slice of `n_items` owned by `rank`: sizes differ by at most one, earlier ranks
    take the extra item, slices tile [0, n_items) in rank order."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(int(n_items), world)
    lo = rank * base + min(rank, extra)
    return slice(lo, lo + base + (1 if rank < extra else 0))


class ShardedBatchSampler:
    """Batch sampler for `torch.utils.data.DataLoader(batch_sampler=...)`.

    `sampler` is any index iterable (a `WeightedRandomSampler` built from the reference's `*_weight.csv`,
    run.py:84-85, a `RandomSampler`, or `range(len(ds))`) that yields the SAME sequence on every rank (seed
    its generator identically).  Each global batch of `world * batch_size` indices is cut into `world`
    contiguous slices; rank r yields slice r.  The union over ranks of step k is therefore exactly the k-th
    batch a single process with batch size `world * batch_size` would have seen.

    drop_last=False: a short final global batch is split with `rank_slice` (some ranks may get an empty
    list and skip the step -- fine for the collective-free frontend; set drop_last=True when a DDP model
    consumes the features and every rank must step the same number of times).
    """

    def __init__(self, sampler: Iterable[int], batch_size: int, rank: Optional[int] = None, world: Optional[int] = None,
                 drop_last: bool = False):
        env_rank, env_world, _ = dist_env()
        self.sampler = sampler
        self.batch_size = int(batch_size)
        self.rank = env_rank if rank is None else int(rank)
        self.world = env_world if world is None else int(world)
        self.drop_last = bool(drop_last)
        if self.batch_size < 1 or not (0 <= self.rank < self.world):
            raise ValueError("bad batch_size / rank / world")

    def __iter__(self) -> Iterator[List[int]]:
        g = self.world * self.batch_size
        buf: List[int] = []
        for idx in self.sampler:
            buf.append(int(idx))
            if len(buf) == g:
                yield buf[self.rank * self.batch_size:(self.rank + 1) * self.batch_size]
                buf = []
        if buf and not self.drop_last:
            mine = buf[rank_slice(len(buf), self.rank, self.world)]
            if mine:
                yield mine

    def __len__(self) -> int:
        n = len(self.sampler)  # type: ignore[arg-type]
        g = self.world * self.batch_size
        full, rem = divmod(n, g)
        if self.drop_last or rem == 0:
            return full
        return full + (1 if len(range(rem)[rank_slice(rem, self.rank, self.world)]) > 0 else 0)


def shard_clip_ids(clip_ids: Sequence[int], rank: int, world: int) -> List[int]:
    """Slice of an explicit clip list (one-click inference over the chunks of a long recording, BASELINE
    config 4): rank r processes a contiguous run of chunks so that its outputs concatenate in rank order."""
    return list(clip_ids[rank_slice(len(clip_ids), rank, world)])
