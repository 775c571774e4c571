#!/usr/bin/env python
"""Throughput of the other BASELINE.json configs (run under gpurun; writes gpurun_out/sweep.json).

C3  FSD50K shape: ragged clips up to > 30 s padded / truncated to target_length = 3000, mixup across unequal lengths,
    SpecAugment 48 / 192: B in {24, 96, 384}.
C4  hour-long 16 kHz recording chunked into 360 eval clips of 10 s (own mean removal, no masks): batch sweep 1 .. 4096
    (the hour is tiled to reach 4096 clips = 2.6 GB in + 2.2 GB out); clips/s and algorithmic GB/s per batch size --
    latency-bound at small B, bandwidth / issue-bound at large B.
Device-resident fp32 inputs, CUDA events, inputs + outputs of a timed run larger than L2 or rotated over copies.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from psla_b200 import params as P  # noqa: E402
from psla_b200.frontend import GpuFbankFrontend, PackedWaves  # noqa: E402

MEAN, STD = -4.6476, 4.5699
OVERLAP = os.environ.get("PSLA_SWEEP_OVERLAP", "stream")  # launch-overlap mode of the timed loops ("call": the library default)
PEAK = 6533.2
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timed(fn, n_warm, n_iter):
    import time
    t0 = time.perf_counter()
    k = 0
    while k < n_warm or time.perf_counter() - t0 < 0.05:  # >= 50 ms of work first: a GPU fresh out of idle needs several
        fn()                                              # milliseconds to reach its clocks (r02: 48 vs 38 us per C3 step)
        k += 1
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n_iter):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n_iter  # ms


def c4_sweep(dev):
    conf = {"num_mel_bins": 128, "target_length": 1056, "mean": MEAN, "std": STD}
    fe = GpuFbankFrontend(conf, device=dev)
    fe.set_launch_overlap(OVERLAP)
    g = torch.Generator(device=dev).manual_seed(2)
    hour = 0.1 * torch.randn(57_600_000, device=dev, generator=g)  # one hour at 16 kHz
    n = 160_000
    rows = []
    for B in (1, 2, 4, 8, 16, 32, 64, 128, 256, 360, 512, 1024, 2048, 4096):
        # clip i of the batch = chunk (i mod 360) of the hour; rotate the chunk origin so consecutive calls read new data
        n_sets = 4 if B <= 1024 else 1
        outs = [torch.empty(B, 1056, 128, device=dev) for _ in range(n_sets)]
        sets = []
        for s in range(n_sets):
            idx = (np.arange(B, dtype=np.int64) + 97 * s) % 360
            sets.append(PackedWaves(hour, idx * n, np.full(B, n, dtype=np.int64)))
        src1 = np.arange(B, dtype=np.int32)
        state = {"k": 0}

        def step():
            k = state["k"] % n_sets
            fe.forward_packed(sets[k], src1, None, None, None, out=outs[k])
            state["k"] += 1

        iters = max(3, min(300, int(40000 / max(B, 16))))
        ms = timed(step, 3, iters)
        byts = B * (4 * n + 4 * 1056 * 128)
        rows.append({"B": B, "ms": ms, "clips_per_s": B / ms * 1e3, "GBps": byts / ms * 1e-6, "frac_hbm": byts / ms * 1e-6 / PEAK,
                     "iters": iters})
        print("C4", rows[-1], flush=True)
        del outs, sets
    # the 360 chunks are independent clips: the batched result equals the one-clip-at-a-time result
    idx = np.arange(8, dtype=np.int64)
    o8 = fe.forward_packed(PackedWaves(hour, idx * n, np.full(8, n, dtype=np.int64)), np.arange(8, dtype=np.int32), None, None, None)
    o1 = fe.forward_packed(PackedWaves(hour, idx[5:6] * n, np.full(1, n, dtype=np.int64)), np.zeros(1, dtype=np.int32), None, None, None)
    same = bool(torch.equal(o8[5], o1[0]))
    fe.close()
    return {"rows": rows, "batched_equals_single": same}


def c3_sweep(dev):
    conf = {"num_mel_bins": 128, "target_length": 3000, "mean": MEAN, "std": STD}
    fe = GpuFbankFrontend(conf, device=dev)
    fe.set_launch_overlap(OVERLAP)  # consecutive calls on one stream, nothing in between
    rows = []
    for B in (24, 96, 384):
        rng = np.random.default_rng(1)
        lens = rng.integers(4800, 480_000, B).astype(np.int64)
        lens[0], lens[1], lens[2] = 480_000, 500_000, 130_000  # exactly 30 s, > 30 s (truncated), < 10 s (padded)
        starts = np.concatenate([[0], np.cumsum((lens + 3) // 4 * 4)[:-1]]).astype(np.int64)
        total = int(starts[-1] + lens[-1] + 8)
        g = torch.Generator(device=dev).manual_seed(3)
        sets = [PackedWaves(0.1 * torch.randn(total, device=dev, generator=g), starts, lens) for _ in range(3)]
        outs = [torch.empty(B, 3000, 128, device=dev) for _ in range(3)]
        src1 = np.arange(B, dtype=np.int32)
        src2 = np.where(np.arange(B) % 2 == 0, (np.arange(B) * 7 + 3) % B, -1).astype(np.int32)  # unequal lengths: pad and cut
        lam = np.where(src2 >= 0, rng.beta(10, 10, B), 0.0)
        torch.manual_seed(3)
        masks = np.array([P.draw_masks(48, 192, 128, 3000) for _ in range(B)], dtype=np.int32)
        state = {"k": 0}

        def step():
            k = state["k"] % 3
            fe.forward_packed(sets[k], src1, src2, lam, masks, out=outs[k])
            state["k"] += 1

        ms = min(timed(step, 3, 200) for _ in range(3))
        n2 = np.where(src2 >= 0, np.minimum(lens[np.maximum(src2, 0)], lens), 0)  # partner samples that are read
        byts = int(4 * (lens.sum() + n2.sum()) + B * 4 * 3000 * 128)
        rows.append({"B": B, "ms": ms, "clips_per_s": B / ms * 1e3, "GBps": byts / ms * 1e-6, "frac_hbm": byts / ms * 1e-6 / PEAK,
                     "audio_seconds_per_s": float(lens.sum()) / 16000 / ms * 1e3, "launches_per_step": 2})
        print("C3", rows[-1], flush=True)
    fe.close()
    return {"rows": rows}


def main():
    dev = torch.device("cuda", 0)
    res = {"device": torch.cuda.get_device_name(0), "hbm_peak_gbs": PEAK, "launch_overlap": OVERLAP, "c3_fsd50k": c3_sweep(dev)}
    if not os.environ.get("PSLA_SWEEP_C3_ONLY"):
        res["c4_hour"] = c4_sweep(dev)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "sweep.json"), "w") as f:
        json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
