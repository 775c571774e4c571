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

    import time
    for on in (True, False, True):
        fe.set_meta_stream(on)
        us = time_steps()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for k in range(200):
            fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print(f"descriptor copy on the copy stream: {on}: {us:.2f} us/step; host time per call {(t1 - t0) / 200 * 1e6:.1f} us")
    for on in (True, False, True):
        fe.set_param_meta(on)
        us = time_steps()
        print(f"descriptors as kernel parameters: {on}: {us:.2f} us/step")
    for on in ("stream", "off", "stream"):
        fe.set_launch_overlap(on)
        us = time_steps()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for k in range(3):
            fe.forward_packed(sets[k % 4], src1, src2, lam, masks, out=outs[k % 4])
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print(f"launch overlap: {on}: {us:.2f} us/step; host time per call (3 calls, queue empty) {(t1 - t0) / 3 * 1e6:.1f} us")
    # overlapped launches must give the same results as plain ones
    fe.set_launch_overlap("off")
    plain = []
    for k in range(4):
        fe.forward_packed(sets[k], src1, src2, lam, masks, out=outs[k])
    torch.cuda.synchronize()
    plain = [o.clone() for o in outs]
    fe.set_launch_overlap("stream")
    for rep in range(5):
        for k in range(4):
            outs[k].zero_()
        for k in range(4):
            fe.forward_packed(sets[k], src1, src2, lam, masks, out=outs[k])
        torch.cuda.synchronize()
        print("overlapped launches identical to plain:", all(bool(torch.equal(a, c)) for a, c in zip(plain, outs)))
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
        tr.zero_()
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
        # events -> steps with phase markers
        ev = r[:, 4:]
        phases = {1: {}, 2: {}, 3: {}}
        names = {4: "clip entered", 5: "window staged", 6: "pass 1 done", 7: "pass 2 + power done", 8: "mel done", 9: "epilogue done", 10: "cold loads issued", 11: "prefetch issued", 99: "step end"}
        for wi in np.nonzero(ok)[0]:
            row = ev[wi]
            cyc = row & ((1 << 40) - 1)
            kind = (row >> 40) & 0xF
            n_ev = int((kind > 0).sum())
            opens = [i for i in range(n_ev) if kind[i] <= 3]
            for j, i in enumerate(opens):
                last_open = j + 1 == len(opens)
                if last_open and n_ev == len(row):
                    continue  # record full: the end of this step is unknown
                end_i = opens[j + 1] if not last_open else n_ev
                t_end = int(cyc[end_i]) if not last_open else int(tot[wi])
                prev = int(cyc[i])
                d = phases[int(kind[i])]
                for q in range(i + 1, end_i):
                    d.setdefault(int(kind[q]), []).append(int(cyc[q]) - prev)
                    prev = int(cyc[q])
                d.setdefault(99, []).append(t_end - prev)
                d.setdefault(100, []).append(t_end - int(cyc[i]))
        for kk, name in ((1, "live step"), (2, "cold-started live step"), (3, "dead run")):
            d = phases[kk]
            if not d:
                continue
            print(f"{name}: n={len(d[100])}  whole step cycles", pct(np.array(d[100])))
            for q in [k for k in (4, 10, 5, 6, 7, 11, 8, 9, 99) if k in d]:
                print(f"    -> {names[q]:22s}", pct(np.array(d[q])), f" n={len(d[q])}")
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
