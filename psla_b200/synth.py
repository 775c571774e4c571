"""Seeded synthetic waveforms: broadband, fp32 in [-1, 1], 16 kHz mono.

Used by the tests, the golden-vector generator and bench.py ("data": "synthetic").  The
reference has no network-free data of its own; BASELINE.json asks for "identical synthetic
waveforms" on both sides of every comparison, so they are generated from torch CPU generators
(bit-reproducible across hosts) rather than stored.
"""
from __future__ import annotations

import math
from typing import List

import torch

SAMPLE_RATE = 16000


def synth_wave(seed: int, n: int, kind: str = "gauss", amp: float = 0.1) -> torch.Tensor:
    """[1, n] float32.  kind: 'gauss' (amp*randn), 'am' (4 Hz amplitude-modulated noise),
    'int16' (noise quantised to the int16 PCM grid, what a decoded wav file holds)."""
    g = torch.Generator().manual_seed(int(seed))
    x = amp * torch.randn(1, n, generator=g)
    if kind == "gauss":
        return x
    if kind == "am":
        t = torch.arange(n, dtype=torch.float32) / SAMPLE_RATE
        return x * (0.55 + 0.45 * torch.sin(2.0 * math.pi * 4.0 * t)).unsqueeze(0)
    if kind == "int16":
        return torch.clamp(torch.round(x * 32768.0), -32768, 32767) / 32768.0
    raise ValueError(kind)


KINDS = ("gauss", "am", "int16")

# tiny in-memory dataset shared by oracle/gen_golden.py and the tests
TINY_LENGTHS = (4800, 8000, 16000, 12345, 24000, 40000, 6400, 16000)
TINY_CLASSES = tuple(f"/m/c{i}" for i in range(10))
TINY_LABELS = ("/m/c0", "/m/c1,/m/c2", "/m/c3", "/m/c0,/m/c4", "/m/c5", "/m/c6,/m/c7,/m/c8", "/m/c9", "/m/c1")


def tiny_dataset_waves() -> List[torch.Tensor]:
    return [synth_wave(100 + i, n, KINDS[i % 3]) for i, n in enumerate(TINY_LENGTHS)]


def write_wav16(path: str, wave: torch.Tensor, sample_rate: int = SAMPLE_RATE) -> None:
    """[1, n] / [n] float32 in [-1, 1) -> mono 16-bit PCM wav file (round to nearest, saturate): what the recipes'
    resampled 16 kHz files are (egs/*/prep_*.py)."""
    import wave as _wave

    pcm = torch.clamp(torch.round(wave.reshape(-1).to(torch.float32) * 32768.0), -32768, 32767).to(torch.int16)
    with _wave.open(path, "wb") as f:
        f.setnchannels(1)
        f.setsampwidth(2)
        f.setframerate(sample_rate)
        f.writeframes(pcm.numpy().astype("<i2").tobytes())


def write_tiny_wav_dataset(directory: str, copies: int = 1):
    """The tiny dataset as real files: 8 * copies 16-bit wav files + manifest json + label csv in `directory`.
    Returns (json_path, csv_path).  Shared by oracle/gen_golden.py (reference side) and the dataloader tests."""
    import json
    import os

    waves = tiny_dataset_waves()
    data = []
    for c in range(copies):
        for i, w in enumerate(waves):
            path = os.path.join(directory, f"clip_{c}_{i}.wav")
            write_wav16(path, w)
            data.append({"wav": path, "labels": TINY_LABELS[i]})
    jpath, cpath = os.path.join(directory, "tiny_files.json"), os.path.join(directory, "labels.csv")
    with open(jpath, "w") as f:
        json.dump({"data": data}, f)
    with open(cpath, "w") as f:
        f.write("index,mid,display_name\n")
        for i, mid in enumerate(TINY_CLASSES):
            f.write(f"{i},{mid},class{i}\n")
    return jpath, cpath
