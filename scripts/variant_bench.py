#!/usr/bin/env python
"""Time one build of libpsla_frontend (PSLA_FRONTEND_LIB) on the bench batch and on a steady-state batch.

    PSLA_FRONTEND_LIB=psla_b200/csrc/variants/libpsla_x.so python scripts/variant_bench.py [tag]

Prints one JSON line: us per 100-clip launch with overlapping launches (median / min of `--repeats` timed regions of
`--steps` launches), serialised launches, the 4096-eval-clip steady state, and a checksum (what-if builds differ).
Used by scripts/gpu_variants.sh to compare kernel experiments inside ONE gpurun call.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bench import B_CLIPS, N_MEL, N_SAMPLES, N_WAV, NORM_MEAN, NORM_STD, T_FRAMES  # noqa: E402
from psla_b200 import params as P  # noqa: E402
from psla_b200.frontend import GpuFbankFrontend, PackedWaves  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("tag", nargs="?", default=os.path.basename(os.environ.get("PSLA_FRONTEND_LIB", "default")))
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--repeats", type=int, default=15)
    ap.add_argument("--no-steady", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    conf = {"num_mel_bins": N_MEL, "target_length": T_FRAMES, "mean": NORM_MEAN, "std": NORM_STD}
    gen = torch.Generator(device=dev).manual_seed(1234)
    offsets = np.arange(N_WAV, dtype=np.int64) * N_SAMPLES
    lengths = np.full(N_WAV, N_SAMPLES, dtype=np.int64)
    sets = [PackedWaves(0.1 * torch.randn(N_WAV * N_SAMPLES, device=dev, generator=gen), offsets, lengths) for _ in range(4)]
    outs = [torch.empty(B_CLIPS, T_FRAMES, N_MEL, device=dev) for _ in range(4)]
    b = np.arange(B_CLIPS)
    src1 = b.astype(np.int32)
    src2 = np.where(b % 2 == 0, B_CLIPS + b // 2, -1).astype(np.int32)
    lam = np.where(b % 2 == 0, np.random.default_rng(0).beta(10, 10, B_CLIPS), 0.0)
    torch.manual_seed(0)
    masks = np.array([P.draw_masks(48, 192, N_MEL, T_FRAMES) for _ in range(B_CLIPS)], dtype=np.int32)
    fe = GpuFbankFrontend(conf, device=dev)

    def region(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(n):
            fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / n

    res = {"tag": a.tag}
    for mode, on in (("overlap", "stream"), ("serial", "off")):
        fe.set_launch_overlap(on)
        region(50)
        us = [region(a.steps) for _ in range(a.repeats)]
        res[mode] = {"median_us": round(statistics.median(us), 3), "min_us": round(min(us), 3), "max_us": round(max(us), 3)}
    fe.set_launch_overlap("stream")
    res["checksum"] = float(outs[0].double().sum().item())
    # eval clips only (no mixup, no masks), same batch size
    region_eval = []
    for _ in range(a.repeats):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(a.steps):
            fe.forward_packed(sets[k % 4], src1, None, None, None, out=outs[k % 4])
        e1.record()
        torch.cuda.synchronize()
        region_eval.append(e0.elapsed_time(e1) * 1e3 / a.steps)
    res["eval100"] = {"median_us": round(statistics.median(region_eval), 3), "min_us": round(min(region_eval), 3)}
    if not a.no_steady:
        hour = 0.1 * torch.randn(57_600_000, device=dev)
        B = 4096
        idx = np.arange(B, dtype=np.int64) % 360
        pk = PackedWaves(hour, idx * N_SAMPLES, np.full(B, N_SAMPLES, dtype=np.int64))
        out = torch.empty(B, T_FRAMES, N_MEL, device=dev)
        s1 = np.arange(B, dtype=np.int32)
        fe.set_param_meta(False)  # ONE launch over all 4096 clips: no launch effects
        for _ in range(2):
            fe.forward_packed(pk, s1, None, None, None, out=out)
        torch.cuda.synchronize()
        t = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fe.forward_packed(pk, s1, None, None, None, out=out)
            e1.record()
            torch.cuda.synchronize()
            t.append(e0.elapsed_time(e1) * 1e3)
        res["steady4096"] = {"us_per_100_clips": round(statistics.median(t) * 100 / B, 3)}
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
