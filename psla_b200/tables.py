"""Host-side constant tables built with torch CPU ops in torchaudio's fp32 op order.

libpsla_frontend.so builds the same tables natively (csrc/tables.cc) with libm's logf/cosf; torch's
vectorised log differs from libm by an ulp on a couple of entries, which moves two mel weights by
~1e-6.  To be bit-faithful to what `kaldi.fbank` multiplies by on this host, the Python frontend
installs these tables through psla_frontend_set_tables().  One-time init, not on the hot path.
"""
from __future__ import annotations

import math

import torch


def hann_window(size: int = 400) -> torch.Tensor:
    """torchaudio/compliance/kaldi.py:94-95 (window_type='hanning')."""
    return torch.hann_window(size, periodic=False, dtype=torch.float32)


def mel_banks(num_bins: int = 128, padded: int = 512, sample_rate: float = 16000.0, low_freq: float = 20.0,
              high_freq: float = 0.0) -> torch.Tensor:
    """Dense [num_bins, padded/2] HTK-mel triangle weights, op order of kaldi.py:436-511 (no VTLN)."""
    nyquist = 0.5 * sample_rate
    if high_freq <= 0.0:
        high_freq += nyquist
    bin_width = sample_rate / padded
    lo = 1127.0 * math.log(1.0 + low_freq / 700.0)
    hi = 1127.0 * math.log(1.0 + high_freq / 700.0)
    step = (hi - lo) / (num_bins + 1)
    idx = torch.arange(num_bins).unsqueeze(1)
    left, center, right = lo + idx * step, lo + (idx + 1.0) * step, lo + (idx + 2.0) * step
    mel = (1127.0 * (1.0 + (bin_width * torch.arange(padded / 2)) / 700.0).log()).unsqueeze(0)
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    return torch.max(torch.zeros(1), torch.min(up, down)).contiguous()
