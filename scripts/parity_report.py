#!/usr/bin/env python
"""Parity numbers as a REPORT (profiles/r02*/parity.json), not only assert bounds: per waveform kind the number of cells whose
log-mel value differs by more than 1e-3 and the largest difference, for
    ours (CUDA, default and exact-clip-mean mode)  vs  the fp32 reference path (oracle: torchaudio kaldi.fbank),
    the fp32 reference path                        vs  an fp64 evaluation of the same recipe,
    ours                                           vs  that fp64 truth.
Kinds: the three synthetic families of the tests (gauss / am / int16), plus adversarial ones for the default mode's
"subtract the clip's first sample" constant: a full-scale click at sample 0 in front of a quiet tonal or
low-pass body, and a loud onset followed by a quiet body.  TEST INFRASTRUCTURE (imports oracle/).

    python scripts/parity_report.py [out.json]
"""
from __future__ import annotations

import json
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import psla_oracle as O  # noqa: E402
from psla_b200.frontend import GpuFbankFrontend  # noqa: E402
from psla_b200.synth import synth_wave  # noqa: E402

MEAN, STD, T = -4.6476, 4.5699, 1056


def tonal(seed, n, amp):
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(n, dtype=torch.float64) / 16000.0
    x = torch.zeros(n, dtype=torch.float64)
    for f in (220.0, 447.3, 1318.5, 2093.0):
        x += torch.sin(2 * math.pi * f * t + float(torch.rand(1, generator=g)) * 6.28)
    return (amp * x / 4).float().unsqueeze(0)


def lowpass(seed, n, amp):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n + 64, generator=g, dtype=torch.float64)
    k = torch.hann_window(65, periodic=False, dtype=torch.float64)
    y = torch.nn.functional.conv1d(x.view(1, 1, -1), (k / k.sum()).view(1, 1, -1)).view(-1)[:n]
    return (amp * y / y.std()).float().unsqueeze(0)


def make(kind, seed, n=160000):
    if kind in ("gauss", "am", "int16"):
        return synth_wave(seed, n, kind)
    if kind == "click_tonal":
        x = tonal(seed, n, 1e-3)
        x[0, 0] = 0.9
        return x
    if kind == "click_lowpass":
        x = lowpass(seed, n, 1e-3)
        x[0, 0] = -0.9
        return x
    if kind == "click_lowpass_int16":
        x = lowpass(seed, n, 4e-3)
        x[0, 0] = 0.9
        return torch.clamp(torch.round(x * 32768.0), -32768, 32767) / 32768.0
    if kind == "loud_onset":
        x = lowpass(seed, n, 1e-3)
        x[0, :2000] += 0.5 * synth_wave(seed + 1, 2000, "gauss", amp=1.0)[0].clamp(-1, 1)
        return x
    if kind == "dc_offset":
        return synth_wave(seed, n, "gauss", amp=1e-3) + 0.5
    raise ValueError(kind)


def stats(a, b, live):
    d = ((a.double() - b.double()).abs() * STD)[live]
    return {"cells": int(live.sum()), "gt_1e-3": int((d > 1e-3).sum()), "max": float(d.max()), "rms": float(d.pow(2).mean().sqrt())}


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "parity.json")
    dev = torch.device("cuda", 0)
    conf = {"num_mel_bins": 128, "target_length": T, "mean": MEAN, "std": STD}
    fes = {"default": GpuFbankFrontend(conf, device=dev), "exact_clip_mean": GpuFbankFrontend(conf, device=dev, exact_clip_mean=True)}
    torch.set_num_threads(max(1, min(16, os.cpu_count() or 1)))
    fill = ((torch.zeros(1) - MEAN) / STD).item()
    report = {"tolerance": 1e-3, "units": "log-mel (normalised outputs multiplied back by std)", "clips_per_kind": 6, "kinds": {}}
    for kind in ("gauss", "am", "int16", "click_tonal", "click_lowpass", "click_lowpass_int16", "loud_onset", "dc_offset"):
        waves = [make(kind, 1000 + s) for s in range(6)]
        B = len(waves)
        src1, src2 = list(range(B)), [1, -1, 3, -1, -1, 0]  # two mixed clips per kind
        lam = [0.37, 0.0, 0.61, 0.0, 0.0, 0.5]
        masks = [(0, 0, 0, 0)] * B
        ref32 = O.batch_torch(waves, src1, src2, lam, masks, T, MEAN, STD)
        ref64 = O.batch_torch(waves, src1, src2, lam, masks, T, MEAN, STD, dtype=torch.float64)
        live = ref32 != fill
        rec = {"ref32_vs_fp64": stats(ref32, ref64, live)}
        for name, fe in fes.items():
            got = fe(waves, src1, src2, lam, masks).cpu()
            assert torch.equal(got == fill, ref32 == fill) or int(((got == fill) ^ (ref32 == fill)).sum()) <= 2
            rec[f"ours_{name}_vs_ref32"] = stats(got, ref32, live)
            rec[f"ours_{name}_vs_fp64"] = stats(got, ref64, live)
        report["kinds"][kind] = rec
        print(kind, json.dumps(rec), flush=True)
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    with open(out_path, "w") as f:
        json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
