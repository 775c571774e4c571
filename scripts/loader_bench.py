#!/usr/bin/env python
"""Clips/s of the WORKER-FED frontend: 16-bit wav files on tmpfs -> torch DataLoader worker processes (decode to int16 PCM,
mixup partner, SpecAugment draws, pack into page-locked shared slots) -> side-stream H2D -> fused kernel -> features on
the GPU.  This is the figure that corresponds to the reference's `per_sample_data_time` (src/traintest.py:87-88): the
frontend fed the way the reference's training loop feeds it (src/run.py:82-94), with the AudioSet training configuration.

    python scripts/loader_bench.py [--seconds 4] [--workers N] [--files 300]

bench.py calls run() for its `e2e.dataloader` sub-record.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_files(directory: str, n_files: int, seconds: float = 10.0, seed: int = 0):
    """n_files mono 16-bit 16 kHz wav files of `seconds` + manifest + label csv (527 classes like AudioSet)."""
    import wave

    n = int(seconds * 16000)
    g = torch.Generator().manual_seed(seed)
    base = (0.1 * torch.randn(n + 4096 * 8, generator=g) * 32768.0).round().clamp(-32768, 32767).to(torch.int16).numpy()
    data = []
    for i in range(n_files):
        path = os.path.join(directory, f"c{i}.wav")
        off = (i * 997) % (4096 * 8)
        with wave.open(path, "wb") as f:
            f.setnchannels(1)
            f.setsampwidth(2)
            f.setframerate(16000)
            f.writeframes(base[off:off + n].astype("<i2").tobytes())
        data.append({"wav": path, "labels": f"/m/c{i % 527},/m/c{(7 * i + 3) % 527}"})
    jpath, cpath = os.path.join(directory, "files.json"), os.path.join(directory, "labels.csv")
    with open(jpath, "w") as f:
        json.dump({"data": data}, f)
    with open(cpath, "w") as f:
        f.write("index,mid,display_name\n")
        for i in range(527):
            f.write(f"{i},/m/c{i},class{i}\n")
    return jpath, cpath


def run(device, seconds: float = 4.0, workers: int | None = None, n_files: int = 300, batch: int = 100, rank: int = 0) -> dict:
    from psla_b200.dataloaders import AudiosetDataset, FrontendLoader, pcm16_lazy_loader
    from psla_b200.frontend import GpuFbankFrontend

    cores = len(os.sched_getaffinity(0))
    workers = max(1, cores - 1) if workers is None else int(workers)
    conf = {"num_mel_bins": 128, "target_length": 1056, "freqm": 48, "timem": 192, "mixup": 0.5, "dataset": "audioset",
            "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=base, prefix=f"psla_lb{rank}_") as d:
        jpath, cpath = make_files(d, n_files, seed=rank)
        ds = AudiosetDataset(jpath, conf, label_csv=cpath, loader=pcm16_lazy_loader, verbose=False)
        fe = GpuFbankFrontend(conf, device=device)
        sampler = torch.utils.data.RandomSampler(ds, replacement=True, num_samples=10 ** 9)  # an endless weighted-style draw
        fl = FrontendLoader.create(ds, fe, batch_size=batch, num_workers=workers, sampler=sampler, max_clip_seconds=10.0,
                                   prefetch_factor=2, persistent_workers=False)
        it = iter(fl)
        acc = torch.zeros((), device=device, dtype=torch.float64)
        for _ in range(max(3, workers // 2)):  # warm-up: workers started, pool cycling
            fb, lab = next(it)
        torch.cuda.synchronize(device)
        l0 = fe.launch_count
        t0 = time.perf_counter()
        steps = clips = 0
        while time.perf_counter() - t0 < seconds:
            fb, lab = next(it)
            acc += fb[:, 0, 0].double().sum()  # a device-side consumer of every batch
            steps += 1
            clips += fb.shape[0]
        checksum = float(acc.item())  # device -> host read: everything above has finished
        t1 = time.perf_counter()
        launches = fe.launch_count - l0
        it.close()
        fl.close()
        per_sample_us = fl.per_sample_data_time.avg * 1e6
        del it, fl
        res = {"value": clips / (t1 - t0), "unit": "clips/s", "steps": steps, "seconds": t1 - t0, "workers": workers, "host_cores": cores,
               "clips_per_step": batch, "gpu_launches": launches, "data_time_per_sample_us": per_sample_us,
               "input": f"{n_files} x 10 s 16-bit wav files on tmpfs, drawn with replacement; AudioSet train conf (mixup 0.5, freqm 48, "
                        "timem 192); workers decode to int16 PCM and pack into page-locked shared slots; features stay on the GPU",
               "checksum": checksum}
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=4.0)
    ap.add_argument("--workers", type=int, default=None)
    ap.add_argument("--files", type=int, default=300)
    a = ap.parse_args()
    print(json.dumps(run(torch.device("cuda", 0), a.seconds, a.workers, a.files)))
