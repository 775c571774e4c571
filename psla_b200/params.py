"""Host-drawn random parameters of the frontend, consumed in the reference's order.

`AudiosetDataset.__getitem__` (src/dataloaders/audioset_dataset.py:139-202) draws, per item:
  random.random()            mixup decision                     (:147)
  random.randint(0, N-1)     mixup partner                      (:153)
  np.random.beta(10, 10)     mix lambda                         (:114)
  torch.rand(1) x2           frequency mask, only if freqm != 0 (:183 -> functional.py:939-940)
  torch.rand(1) x2           time mask, only if timem != 0      (:185)
  torch.rand(T, F), np.random.rand(), np.random.randint(-10, 10)   noise field, its scale, roll -- only if noise (:197-199)
The GPU kernels take the resulting integers / lambda as inputs, so with the same seeds the host
reproduces the reference's augmentation decisions exactly.
"""
from __future__ import annotations

import random as _py_random
from dataclasses import dataclass
from typing import Tuple

import numpy as np
import torch


@dataclass
class ItemParams:
    mixed: bool
    partner: int  # dataset index of the partner, -1 if not mixed
    lam: float  # float64 lambda, 0.0 if not mixed (what _wav2fbank returns, :134-137)
    mask: Tuple[int, int, int, int]  # f_start, f_width, t_start, t_width (width 0 = no mask)
    noise_scale: float = 0.0  # np.random.rand() of :198 (0.0 when the noise augmentation is off)
    roll: int = 0  # np.random.randint(-10, 10) of :199
    noise: bool = False  # the noise / roll augmentation applies to this item


def axis_mask(mask_param: int, axis_size: int) -> Tuple[int, int]:
    """(start, width) exactly as torchaudio.functional.mask_along_axis computes them in fp32
    (functional.py:939-943): value = rand*P ; min_value = rand*(size - value) ; long() truncation."""
    # the same two draws from torch's global generator, then the same IEEE fp32 products on numpy scalars (a dataloader
    # worker spends more time in five tiny tensor ops than in decoding the file): bit-identical to the tensor expression,
    # checked against the real FrequencyMasking / TimeMasking modules over 200 seeds (tests/golden/mask_offsets.npz)
    r1, r2 = np.float32(torch.rand(1).item()), np.float32(torch.rand(1).item())
    value = r1 * np.float32(mask_param)
    min_value = r2 * (np.float32(axis_size) - value)
    return int(min_value), int(value)


def draw_masks(freqm: int, timem: int, num_mel_bins: int, target_length: int) -> Tuple[int, int, int, int]:
    f0 = fw = t0 = tw = 0
    if freqm != 0:
        f0, fw = axis_mask(freqm, num_mel_bins)
    if timem != 0:
        t0, tw = axis_mask(timem, target_length)
    return f0, fw, t0, tw


def draw_mixup(mixup: float, n_items: int, py_random=_py_random) -> Tuple[bool, int]:
    if py_random.random() < mixup:
        return True, py_random.randint(0, n_items - 1)
    return False, -1


def draw_lambda(np_random=np.random) -> float:
    return float(np_random.beta(10, 10))


def draw_noise(target_length: int, num_mel_bins: int, np_random=np.random) -> Tuple[torch.Tensor, float, int]:
    """The three draws of the noise / roll augmentation in the reference's order (:198-199): the U[0,1) field from
    torch's generator, then the scale and the roll from numpy's."""
    field = torch.rand(target_length, num_mel_bins)
    scale = float(np_random.rand())
    roll = int(np_random.randint(-10, 10))
    return field, scale, roll
