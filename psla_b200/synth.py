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
