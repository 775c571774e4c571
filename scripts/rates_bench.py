#!/usr/bin/env python
"""Throughput of the geometry-generic kernel (sample rates other than 16 kHz) next to the streaming kernel.

    python scripts/rates_bench.py [out.json]

Per sample rate: B eval clips of 10 s (fp32, device-resident, T = the clip's own frame count), CUDA events over back-to-back
launches after a warm-up of >= 50 ms, clips/s, audio seconds per second, and the HBM fraction of SURVEY 8(d)'s byte count
(4 bytes per input sample + 4 bytes per output cell) against MEASURED_PEAKS.json.  16 kHz is listed twice: streaming kernel
(exact clip means, i.e. with the statistics pass the generic path always runs) and the generic kernel.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from psla_b200.frontend import GpuFbankFrontend, PackedWaves  # noqa: E402


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6550.0


def run(fe, sr, B, dev, iters=20):
    n = 10 * sr
    T = fe.num_frames(n)
    fe.target_length = T
    wav = 0.1 * torch.randn(B * n, device=dev)
    pk = PackedWaves(wav, np.arange(B, dtype=np.int64) * n, np.full(B, n, dtype=np.int64))
    src1 = np.arange(B, dtype=np.int32)
    outs = [torch.empty(B, T, fe.num_mel_bins, device=dev) for _ in range(2)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t = 0.0
    while t < 60.0:  # warm-up: clocks
        e0.record()
        fe.forward_packed(pk, src1, None, None, None, out=outs[0], target_length=T)
        e1.record()
        torch.cuda.synchronize()
        t += e0.elapsed_time(e1)
    best = None
    for _ in range(3):
        e0.record()
        for k in range(iters):
            fe.forward_packed(pk, src1, None, None, None, out=outs[k % 2], target_length=T)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        best = ms if best is None else min(best, ms)
    bytes_ = B * (4 * n + 4 * T * fe.num_mel_bins)
    return {"sample_rate": sr, "frame": [fe.frame_length, fe.frame_shift, fe.fft_size], "generic_path": fe.generic_path, "B": B, "T": T,
            "ms": best, "clips_per_s": B / best * 1e3, "audio_seconds_per_s": 10 * B / best * 1e3,
            "GBps": bytes_ / best / 1e6, "frac_hbm": bytes_ / best / 1e6 / peak_gbs()}


def main():
    dev = torch.device("cuda", 0)
    rows = []
    B = 100
    for sr in (8000, 16000, 22050, 32000, 44100, 48000):
        conf = {"num_mel_bins": 128, "target_length": 1056, "mean": -4.6476, "std": 4.5699}
        fe = GpuFbankFrontend(conf, device=dev, sample_rate=sr, exact_clip_mean=True)
        rows.append(run(fe, sr, B, dev))
        print(json.dumps(rows[-1]), flush=True)
        if sr == 16000:
            fe.set_generic_path(True)
            rows.append(run(fe, sr, B, dev))
            print(json.dumps(rows[-1]), flush=True)
        fe.close()
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "rates_bench.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(rows, open(out, "w"), indent=1)


if __name__ == "__main__":
    main()
