#!/usr/bin/env python
"""Per-warp timeline of the streaming kernel on the bench workload (psla_frontend_set_trace).

Prints where a launch's time goes: the spread of warp start / end times (globaltimer), the cycles a warp spends before
its first group (table copy + partition search), cycles per live group, and the imbalance between warps and SMs.
Writes the raw records to gpurun_out/trace.npz.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from bench import B_CLIPS, N_MEL, N_SAMPLES, N_WAV, NORM_MEAN, NORM_STD, T_FRAMES  # noqa: E402
from psla_b200 import params as P  # noqa: E402
from psla_b200.frontend import GpuFbankFrontend, PackedWaves  # noqa: E402


def pct(x, qs=(0, 5, 25, 50, 75, 95, 100)):
    return " ".join(f"{np.percentile(x, q):9.1f}" for q in qs)


def main():
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
    for k in range(20):
        fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
    torch.cuda.synchronize()
    def time_steps(n=200):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for k in range(10):
            fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
        e0.record()
        for k in range(n):
            fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / n

    ref = outs[0].clone()
    for cm in [(64, 2, 48), (64, 2, 0), (64, 4, 48), (64, 1, 48), (64, 2, 32), (64, 2, 64), (64, 8, 48), (32, 1, 0)]:
        fe.set_cost_model(*cm)
        us = time_steps()
        fe.forward_packed(sets[0], src1, src2, lam, masks, out=outs[0])
        torch.cuda.synchronize()
        print(f"cost model live/dead/cold {cm}: {us:.2f} us/step; output identical: {bool(torch.equal(ref, outs[0]))}")
    fe.set_cost_model(64, 2, 48)

    tr = fe.set_trace(True)
    recs = []
    for k in range(8):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
        e1.record()
        torch.cuda.synchronize()
        recs.append((e0.elapsed_time(e1) * 1e3, tr.cpu().numpy().copy()))
    fe.set_trace(False)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    np.savez(os.path.join(ROOT, "gpurun_out", "trace.npz"), us=np.array([r[0] for r in recs]), rec=np.stack([r[1] for r in recs]))
    for us, r in recs[-2:]:
        t0, t1 = r[:, 0], r[:, 1]
        pro = r[:, 2] & 0xFFFFFFFF
        tot = (r[:, 2] >> 32) & 0xFFFFFFFF
        smid = r[:, 3] & 0xFFFF
        live = (r[:, 3] >> 16) & 0xFFFF
        dead = (r[:, 3] >> 32) & 0xFFFF
        cold = (r[:, 3] >> 48) & 0xFFFF
        ok = t0 > 0
        base = t0[ok].min()
        print(f"--- launch: events {us:.1f} us; warps {ok.sum()} ; SMs {len(set(smid[ok].tolist()))}; span {(t1[ok].max() - base) / 1e3:.1f} us")
        print("percentiles            min        p5       p25       p50       p75       p95       max")
        print("start  (us)     ", pct((t0[ok] - base) / 1e3))
        print("end    (us)     ", pct((t1[ok] - base) / 1e3))
        print("prologue cycles ", pct(pro[ok]))
        print("total cycles    ", pct(tot[ok]))
        print("live groups     ", pct(live[ok]))
        print("dead groups     ", pct(dead[ok]))
        print("cold starts     ", pct(cold[ok]))
        # per-step durations by kind and by position in the warp's sequence
        steps = r[:, 4:]
        cyc = steps & ((1 << 40) - 1)
        kind = (steps >> 40) & 0xF
        ngrp = (steps >> 44) & 0xFFFF
        nxt = np.concatenate([cyc[:, 1:], np.zeros((len(cyc), 1), dtype=cyc.dtype)], axis=1)
        valid_next = np.concatenate([kind[:, 1:] > 0, np.zeros((len(cyc), 1), dtype=bool)], axis=1)
        last = (kind > 0) & ~valid_next
        dur = np.where(valid_next, nxt - cyc, 0)
        dur = np.where(last, tot[:, None] - cyc, dur)  # the last step ends with the warp
        for kk, name in ((1, "live"), (2, "cold+live"), (3, "dead run")):
            m = (kind == kk) & ok[:, None]
            if m.any():
                d = dur[m]
                print(f"step cycles {name:10s}", pct(d), f"  n={m.sum()}  groups/step {ngrp[m].mean():.1f}")
        for pos in range(0, 12):
            m = (kind[:, pos] == 1) & ok
            if m.any():
                print(f"  live step at position {pos:2d}: median {np.median(dur[m, pos]):8.0f} cycles  start at {np.median(cyc[m, pos]):8.0f}")
        sm_end = {}
        for s_, e, lv in zip(smid[ok], t1[ok], live[ok]):
            a_ = sm_end.setdefault(int(s_), [0, 0])
            a_[0] = max(a_[0], (e - base) / 1e3)
            a_[1] += int(lv)
        ends = np.array([v[0] for v in sm_end.values()])
        print("SM end (us)     ", pct(ends))
        ghz = (tot[ok] / np.maximum(t1[ok] - t0[ok], 1)).mean()
        print(f"mean clock {ghz:.3f} GHz; total live groups {live[ok].sum()}, dead {dead[ok].sum()}")


if __name__ == "__main__":
    main()
