"""Timed CPU baseline for bench.py: the reference's dataloader path, restated (oracle port).  TEST/BENCH ONLY.

`/root/reference` cannot travel to the GPU box and has no installable package (no setup.py), so the reference
arm is the oracle port of `AudiosetDataset.__getitem__` (oracle/psla_oracle.py::item_torch), which calls the same
third-party `torchaudio.compliance.kaldi.fbank` the reference calls at audioset_dataset.py:119-120, driven the way
the reference drives it (src/run.py:87-94): a `torch.utils.data.DataLoader` with `num_workers` processes, one
intra-op thread each, default collate to [B, T, 128].  File decode is replaced by in-memory synthetic waveforms
(the decode is outside the hot path), which flatters the CPU side slightly.
"""
from __future__ import annotations

import os
import random
import sys
import time
from typing import Dict

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from oracle import psla_oracle as O  # noqa: E402
from psla_b200.synth import synth_wave  # noqa: E402

TRAIN_CONF = {"num_mel_bins": 128, "target_length": 1056, "freqm": 48, "timem": 192, "mixup": 0.5,
              "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}


class OracleDataset(Dataset):
    """len() virtual items over a small pool of in-memory 10 s waveforms (shared with forked workers)."""

    def __init__(self, conf: dict, n_items: int, pool: int = 32, n_samples: int = 160000):
        self.conf, self.n_items = conf, n_items
        self.waves = [synth_wave(1000 + i, n_samples) for i in range(pool)]

    def __len__(self):
        return self.n_items

    def __getitem__(self, i):
        fb, _ = O.item_torch(self.waves, i % len(self.waves), self.conf, random, np.random, use_torchaudio=True)
        return fb, torch.zeros(1)


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def _loader(conf, batch, workers, n_items):
    # Each worker produces SUB-batches of `sub` clips so that the 100 clips of one step are spread over all
    # cores (a step = ceil(batch / sub) sub-batches); at most 2 steps of work are prefetched, so timing K steps
    # after a >= 3-step warm-up measures the steady-state arrival rate without a prefetch head start.
    sub = max(1, batch // max(workers, 1))
    ds = OracleDataset(conf, n_items)
    dl = DataLoader(ds, batch_size=sub, shuffle=False, num_workers=workers, pin_memory=False,
                    prefetch_factor=2 if workers > 0 else None)
    return iter(dl), sub


def _step(it, batch):
    n, parts = 0, []
    while n < batch:
        fb, _ = next(it)
        parts.append(fb)
        n += fb.shape[0]
    return n


def time_steps(steps: int, warmup: int, batch: int = 100, workers: int | None = None, conf: dict | None = None,
               seconds: float | None = None) -> Dict:
    """Time `steps` steps (one step = `batch` clips through the DataLoader) after `warmup` (>= 3) untimed steps,
    all host cores as dataloader workers.  With `seconds`, stop early once that much time has been measured."""
    conf = dict(conf or TRAIN_CONF)
    workers = host_cores() if workers is None else workers
    warmup = max(3, warmup)
    it, sub = _loader(conf, batch, workers, (steps + warmup + 4) * (batch + sub_pad(batch, workers)))
    for _ in range(warmup):
        _step(it, batch)
    t0 = time.perf_counter()
    n, done = 0, 0
    for _ in range(steps):
        n += _step(it, batch)
        done += 1
        if seconds is not None and time.perf_counter() - t0 >= seconds:
            break
    dt = time.perf_counter() - t0
    del it
    return {"clips_per_s": n / dt, "seconds": dt, "clips": n, "steps": done, "cores": max(workers, 1),
            "ms_per_step": dt / done * 1e3,
            "sample": f"{done} steps x {batch} clips (10 s @16 kHz, train conf: mixup 0.5, freqm 48, timem 192, T=1056) after "
                      f"{warmup} warm-up steps; torch DataLoader num_workers={workers} (sub-batches of {sub}), 1 thread/worker, "
                      f"in-memory synthetic waveforms, torchaudio kaldi.fbank"}


def sub_pad(batch, workers):
    return max(1, batch // max(workers, 1))


def time_budget(seconds: float = 12.0, batch: int = 100, workers: int | None = None) -> Dict:
    """Bounded sample for bench.py's cpu_baseline leg: about `seconds` of steady-state work."""
    return time_steps(100000, 3, batch, workers, seconds=seconds)


if __name__ == "__main__":
    print(time_budget(float(sys.argv[1]) if len(sys.argv) > 1 else 5.0))
