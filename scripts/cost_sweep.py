#!/usr/bin/env python
"""Sweep of the work-partition cost model (psla_frontend_set_cost_model: live / dead / cold-start cost of a 4-frame group)
on the bench workload (AudioSet training batch, device-resident inputs, back-to-back launches).  Writes
gpurun_out/cost_sweep.json: ms per launch for every (live, dead, cold), best of 3 repetitions of 200 launches."""
from __future__ import annotations

import itertools
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

B, T, NMEL, N, NWAV, NSETS = 100, 1056, 128, 160000, 150, 4


def main():
    from psla_b200 import params as P
    from psla_b200.frontend import GpuFbankFrontend, PackedWaves
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(1234)
    offs, lens = np.arange(NWAV, dtype=np.int64) * N, np.full(NWAV, N, dtype=np.int64)
    sets = [PackedWaves(0.1 * torch.randn(NWAV * N, device=dev, generator=gen), offs, lens) for _ in range(NSETS)]
    outs = [torch.empty(B, T, NMEL, device=dev) for _ in range(NSETS)]
    b = np.arange(B)
    src1, src2 = b.astype(np.int32), np.where(b % 2 == 0, B + b // 2, -1).astype(np.int32)
    lam = np.where(b % 2 == 0, np.random.default_rng(0).beta(10, 10, B), 0.0)
    torch.manual_seed(0)
    masks = np.array([P.draw_masks(48, 192, NMEL, T) for _ in range(B)], dtype=np.int32)
    fe = GpuFbankFrontend({"num_mel_bins": NMEL, "target_length": T, "mean": -4.6476, "std": 4.5699}, device=dev)

    def timed(steps=200):
        for k in range(10):
            fe.forward_packed(sets[k % NSETS], src1, src2, lam, masks, out=outs[k % NSETS])
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(steps):
            fe.forward_packed(sets[k % NSETS], src1, src2, lam, masks, out=outs[k % NSETS])
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    ref = None
    rows = []
    grid = list(itertools.product((64,), (1, 2, 4, 8), (0, 16, 32, 48, 64, 96, 128)))
    for live, dead, cold in grid:
        fe.set_cost_model(live, dead, cold)
        ms = min(timed() for _ in range(3))
        o = outs[0].clone()
        if ref is None:
            ref = o
        same = bool(torch.equal(o, ref))  # the partition must not change a single bit of the result
        rows.append({"live": live, "dead": dead, "cold": cold, "ms": ms, "bit_identical": same})
        print(rows[-1], flush=True)
    rows.sort(key=lambda r: r["ms"])
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "cost_sweep.json"), "w"), indent=1)
    print("best:", rows[:5])


if __name__ == "__main__":
    main()
