"""CPU oracle for the PSLA feature-frontend hot path.  TEST INFRASTRUCTURE ONLY.

This module is a CPU restatement of what the reference computes between "waveform
loaded" and "normalised [target_length, num_mel_bins] tensor returned by
``AudiosetDataset.__getitem__``".  Only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it, and only
as the checker / the timed CPU baseline -- never as (part of) the product path.
``psla_b200`` must not import anything from ``oracle/``.

Where the arithmetic lives
--------------------------
The reference (``/root/reference/src/dataloaders/audioset_dataset.py``) does not
contain the arithmetic itself; it calls the third-party package ``torchaudio``
(pinned ``torchaudio==0.6.0`` / ``torch==1.6.0`` in ``requirements.txt:8-9``, NOT
vendored under ``/root/reference``):

* ``torchaudio.compliance.kaldi.fbank``      (call site ``audioset_dataset.py:119-120``)
* ``torchaudio.transforms.FrequencyMasking`` / ``TimeMasking`` (``:177-185``)

The image has torchaudio 2.11.0 whose ``compliance/kaldi.py`` is pure-Python torch
code implementing the same published Kaldi ``compute-fbank-feats`` recipe.  The
restatement below follows that file (cited as ``TA:kaldi.py:N``) and
``torchaudio/functional/functional.py`` (``TA:functional.py:N``).

Parity pinning
--------------
The reference ships no tests, golden vectors or fixtures for this path, so the
pins are outputs of the reference itself run in the build container:
``oracle/gen_golden.py`` imports the UNMODIFIED ``/root/reference/src/dataloaders``
(with ``torchaudio.load`` monkey-patched, because torchcodec is not installed),
runs ``AudiosetDataset.__getitem__`` and ``kaldi.fbank`` on seeded synthetic
waveforms and commits the results under ``tests/golden/``.
``tests/test_oracle_golden.py`` checks every function below against them.

Three flavours are provided:

* ``*_torch``  -- fp32, written with torch CPU ops in the reference's op order.  This
  reproduces the reference bit-for-bit on the same host (same MKL FFT / sgemm).
* ``*_np``     -- numpy restatement, dtype selectable (float32 or float64).  The
  float64 flavour is the "truth" used for error attribution: the fp32 reference
  itself is up to ~1.5e-3 away from it on rare near-floor cells.
* ``ref_*``    -- thin wrappers that call the installed torchaudio exactly like the
  reference call sites do (used as the live oracle and as the timed CPU baseline).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

# ----------------------------------------------------------------------------------
# Constants of the reference call site (audioset_dataset.py:119-120 + kaldi.fbank
# defaults, TA:kaldi.py:514-541)
# ----------------------------------------------------------------------------------
SAMPLE_RATE = 16000
FRAME_LENGTH_MS = 25.0  # TA:kaldi.py:520
FRAME_SHIFT_MS = 10.0  # audioset_dataset.py:120
PREEMPH = 0.97  # TA:kaldi.py:527
LOW_FREQ = 20.0  # TA:kaldi.py:524
HIGH_FREQ = 0.0  # TA:kaldi.py:522 (0 => Nyquist)
EPS32 = float(np.finfo(np.float32).eps)  # TA:kaldi.py:22


def frame_geometry(sample_rate: float = SAMPLE_RATE) -> Tuple[int, int, int]:
    """(window_shift, window_size, padded_window_size) -- TA:kaldi.py:138-140."""
    shift = int(sample_rate * FRAME_SHIFT_MS * 0.001)
    size = int(sample_rate * FRAME_LENGTH_MS * 0.001)
    padded = 1 if size == 0 else 2 ** (size - 1).bit_length()
    return shift, size, padded


def num_frames(n_samples: int, sample_rate: float = SAMPLE_RATE) -> int:
    """snip_edges=True frame count -- TA:kaldi.py:63-67.  Raises like TA:kaldi.py:142."""
    shift, size, _ = frame_geometry(sample_rate)
    if n_samples < size:
        raise AssertionError(f"choose a window size {size} that is [2, {n_samples}]")
    return 1 + (n_samples - size) // shift


# ----------------------------------------------------------------------------------
# Mel filterbank and window tables
# ----------------------------------------------------------------------------------
def mel_banks_torch(num_bins: int = 128, padded: int = 512, sample_rate: float = SAMPLE_RATE,
                    low_freq: float = LOW_FREQ, high_freq: float = HIGH_FREQ) -> torch.Tensor:
    """fp32 HTK-mel triangular weights [num_bins, padded/2], torch op order of
    TA:kaldi.py:436-511 (vtln_warp == 1.0 branch)."""
    num_fft_bins = padded / 2
    nyquist = 0.5 * sample_rate
    if high_freq <= 0.0:
        high_freq += nyquist
    fft_bin_width = sample_rate / padded
    mel_low = 1127.0 * math.log(1.0 + low_freq / 700.0)  # TA:kaldi.py:326-327
    mel_high = 1127.0 * math.log(1.0 + high_freq / 700.0)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = torch.arange(num_bins).unsqueeze(1)
    left = mel_low + b * delta
    center = mel_low + (b + 1.0) * delta
    right = mel_low + (b + 2.0) * delta
    mel = (1127.0 * (1.0 + (fft_bin_width * torch.arange(num_fft_bins)) / 700.0).log()).unsqueeze(0)
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    return torch.max(torch.zeros(1), torch.min(up, down))


def mel_banks_np(num_bins: int = 128, padded: int = 512, sample_rate: float = SAMPLE_RATE,
                 low_freq: float = LOW_FREQ, high_freq: float = HIGH_FREQ, dtype=np.float64) -> np.ndarray:
    """numpy restatement of TA:kaldi.py:436-511 in ``dtype`` (float64 = 'naive' weights)."""
    num_fft_bins = padded // 2
    nyquist = 0.5 * sample_rate
    if high_freq <= 0.0:
        high_freq += nyquist
    fft_bin_width = sample_rate / padded
    mel_low = 1127.0 * math.log(1.0 + low_freq / 700.0)
    mel_high = 1127.0 * math.log(1.0 + high_freq / 700.0)
    delta = (mel_high - mel_low) / (num_bins + 1)
    b = np.arange(num_bins, dtype=dtype)[:, None]
    left = (mel_low + b * delta).astype(dtype)
    center = (mel_low + (b + 1.0) * delta).astype(dtype)
    right = (mel_low + (b + 2.0) * delta).astype(dtype)
    f = (fft_bin_width * np.arange(num_fft_bins, dtype=dtype)).astype(dtype)
    mel = (1127.0 * np.log((1.0 + f / 700.0).astype(dtype))).astype(dtype)[None, :]
    up = (mel - left) / (center - left)
    down = (right - mel) / (right - center)
    return np.maximum(0.0, np.minimum(up, down)).astype(dtype)


def hann_window_np(size: int = 400, dtype=np.float64) -> np.ndarray:
    """torch.hann_window(size, periodic=False) -- TA:kaldi.py:94-95.

    ATen builds it in the tensor dtype as ``arange(size).mul_(2*pi/(size-1)).cos_().mul_(-0.5).add_(0.5)``
    (aten/src/ATen/native/TensorFactories.cpp, hamming_window with alpha=beta=0.5), so the fp32
    table is NOT the correctly rounded Hann window (up to 1.8e-7 abs / 7e-5 relative near the
    edges).  The float32 flavour restates that op order; float64 is the exact window."""
    if np.dtype(dtype) == np.float32:
        a = np.arange(size, dtype=np.float32) * np.float32(2.0 * math.pi / (size - 1))
        return (np.cos(a).astype(np.float32) * np.float32(-0.5) + np.float32(0.5)).astype(np.float32)
    j = np.arange(size, dtype=np.float64)
    return (0.5 - 0.5 * np.cos(2.0 * np.pi * j / (size - 1))).astype(dtype)


# ----------------------------------------------------------------------------------
# kaldi.fbank restatements
# ----------------------------------------------------------------------------------
def kaldi_fbank_np(wave: np.ndarray, num_mel_bins: int = 128, sample_rate: float = SAMPLE_RATE,
                   dtype=np.float64, mel: Optional[np.ndarray] = None) -> np.ndarray:
    """numpy restatement of ``kaldi.fbank(wave[None], htk_compat=True, use_energy=False,
    window_type='hanning', dither=0.0, frame_shift=10)`` in ``dtype``.

    ``wave`` is the 1-D channel-0 waveform.  ``mel`` (optional, [num_mel_bins, padded/2])
    overrides the weight table; by default the fp32 torch-order table is used (that is what
    the fp64 call of the real function does too: TA:kaldi.py:482,624 build it in fp32).
    Returns [m, num_mel_bins].
    """
    x = np.asarray(wave).astype(dtype)
    assert x.ndim == 1
    shift, size, padded = frame_geometry(sample_rate)
    m = num_frames(x.shape[0], sample_rate)  # TA:kaldi.py:142, 63-67
    idx = (np.arange(m)[:, None] * shift + np.arange(size)[None, :])
    frames = x[idx]  # TA:kaldi.py:82-83 (as_strided view)
    frames = frames - frames.mean(axis=1, keepdims=True, dtype=dtype)  # TA:kaldi.py:183-186
    prev = np.concatenate([frames[:, :1], frames[:, :-1]], axis=1)  # replicate pad, TA:kaldi.py:193-198
    frames = frames - dtype(PREEMPH) * prev
    frames = frames * hann_window_np(size, dtype)[None, :]  # TA:kaldi.py:201-204
    padded_frames = np.zeros((m, padded), dtype=dtype)  # TA:kaldi.py:207-211
    padded_frames[:, :size] = frames
    spec = np.fft.rfft(padded_frames, axis=1)  # TA:kaldi.py:616
    power = (np.abs(spec).astype(dtype)) ** dtype(2.0)  # TA:kaldi.py:616-618
    if mel is None:
        mel = mel_banks_torch(num_mel_bins, padded, sample_rate).numpy()
    melw = np.zeros((num_mel_bins, padded // 2 + 1), dtype=dtype)  # TA:kaldi.py:627
    melw[:, : padded // 2] = mel.astype(dtype)
    e = power @ melw.T  # TA:kaldi.py:630
    return np.log(np.maximum(e, dtype(EPS32))).astype(dtype)  # TA:kaldi.py:633


def kaldi_fbank_torch(wave: torch.Tensor, num_mel_bins: int = 128, sample_rate: float = SAMPLE_RATE) -> torch.Tensor:
    """torch (CPU) restatement in the reference's op order; dtype follows ``wave`` (1-D)."""
    assert wave.dim() == 1
    dtype = wave.dtype
    shift, size, padded = frame_geometry(sample_rate)
    m = num_frames(wave.shape[0], sample_rate)
    frames = wave.as_strided((m, size), (shift * wave.stride(0), wave.stride(0)))
    frames = frames - torch.mean(frames, dim=1).unsqueeze(1)
    off = torch.nn.functional.pad(frames.unsqueeze(0), (1, 0), mode="replicate").squeeze(0)
    frames = frames - PREEMPH * off[:, :-1]
    frames = frames * torch.hann_window(size, periodic=False, dtype=dtype).unsqueeze(0)
    frames = torch.nn.functional.pad(frames.unsqueeze(0), (0, padded - size), mode="constant", value=0).squeeze(0)
    spec = torch.fft.rfft(frames).abs().pow(2.0)
    mel = mel_banks_torch(num_mel_bins, padded, sample_rate).to(dtype)
    mel = torch.nn.functional.pad(mel, (0, 1), mode="constant", value=0)
    e = torch.mm(spec, mel.T)
    return torch.max(e, torch.tensor(EPS32, dtype=dtype)).log()


def ref_fbank(wave: torch.Tensor, num_mel_bins: int = 128, sample_rate: float = SAMPLE_RATE) -> torch.Tensor:
    """The exact third-party call of ``audioset_dataset.py:119-120``; ``wave`` is [1, n]."""
    import torchaudio

    return torchaudio.compliance.kaldi.fbank(
        wave, htk_compat=True, sample_frequency=sample_rate, use_energy=False,
        window_type="hanning", num_mel_bins=num_mel_bins, dither=0.0, frame_shift=10)


# ----------------------------------------------------------------------------------
# _wav2fbank (audioset_dataset.py:88-137)
# ----------------------------------------------------------------------------------
def mix_waveforms_torch(w1: torch.Tensor, w2: Optional[torch.Tensor], lam: float) -> torch.Tensor:
    """Waveform DC removal and optional mixup -- audioset_dataset.py:91-117.
    ``w1``/``w2`` are [1, n]; ``lam`` is the float64 Beta(10,10) draw.  Returns [1, n1]."""
    if w2 is None:
        return w1 - w1.mean()  # :92
    w1 = w1 - w1.mean()  # :98
    w2 = w2 - w2.mean()  # :99
    if w1.shape[1] != w2.shape[1]:
        if w1.shape[1] > w2.shape[1]:
            tmp = torch.zeros(1, w1.shape[1], dtype=w1.dtype)  # :104-106 (pad AFTER mean removal)
            tmp[0, : w2.shape[1]] = w2
            w2 = tmp
        else:
            w2 = w2[0, : w1.shape[1]]  # :109 (1-D, broadcasts)
    mix = lam * w1 + (1 - lam) * w2  # :116  (python/np float64 scalars -> fp32 scalars)
    return mix - mix.mean()  # :117


def pad_or_truncate(fbank: torch.Tensor, target_length: int) -> torch.Tensor:
    """audioset_dataset.py:122-132: zero rows appended or rows cut."""
    p = target_length - fbank.shape[0]
    if p > 0:
        return torch.nn.functional.pad(fbank, (0, 0, 0, p), mode="constant", value=0.0)
    if p < 0:
        return fbank[:target_length, :]
    return fbank


def wav2fbank_torch(w1: torch.Tensor, w2: Optional[torch.Tensor], lam: float, target_length: int,
                    num_mel_bins: int = 128, sample_rate: float = SAMPLE_RATE, use_torchaudio: bool = True) -> torch.Tensor:
    """_wav2fbank minus the file decode; returns [target_length, num_mel_bins]."""
    wave = mix_waveforms_torch(w1, w2, lam)
    if use_torchaudio:
        fb = ref_fbank(wave, num_mel_bins, sample_rate)
    else:
        fb = kaldi_fbank_torch(wave[0], num_mel_bins, sample_rate)
    return pad_or_truncate(fb, target_length)


# ----------------------------------------------------------------------------------
# SpecAugment offsets (TA:functional.py:939-943) and the tail of __getitem__ (:177-192)
# ----------------------------------------------------------------------------------
def mask_offsets_from_uniform(r_value: torch.Tensor, r_start: torch.Tensor, mask_param: int, axis_size: int) -> Tuple[int, int]:
    """(start, width) from the two fp32 uniforms drawn by ``mask_along_axis``:
    value=r*P ; min_value=r'*(size-value) ; start=long(min_value) ; end=start+long(value)."""
    value = r_value * mask_param
    min_value = r_start * (axis_size - value)
    start = int(min_value.long())
    width = int(value.long())
    return start, width


def draw_mask_offsets(freqm: int, timem: int, num_mel_bins: int, target_length: int) -> Tuple[int, int, int, int]:
    """Consume the torch global CPU generator exactly like audioset_dataset.py:182-185
    (freq first, then time; nothing drawn for a zero parameter).  Returns
    (f_start, f_width, t_start, t_width) with width 0 meaning 'no mask'."""
    f0 = fw = t0 = tw = 0
    if freqm != 0:
        rv = torch.rand(1)
        rs = torch.rand(1)
        f0, fw = mask_offsets_from_uniform(rv, rs, freqm, num_mel_bins)
    if timem != 0:
        rv = torch.rand(1)
        rs = torch.rand(1)
        t0, tw = mask_offsets_from_uniform(rv, rs, timem, target_length)
    return f0, fw, t0, tw


def apply_masks_and_norm_torch(fbank: torch.Tensor, mask: Sequence[int], norm_mean: float, norm_std: float,
                               skip_norm: bool = False) -> torch.Tensor:
    """audioset_dataset.py:177-192 given the already-drawn offsets.  ``fbank`` is [T, F]."""
    f0, fw, t0, tw = (int(v) for v in mask)
    fb = fbank.clone()
    if fw > 0:
        fb[:, f0:f0 + fw] = 0.0  # masked_fill(…, 0.0), TA:functional.py:953
    if tw > 0:
        fb[t0:t0 + tw, :] = 0.0
    if not skip_norm:
        fb = (fb - norm_mean) / (norm_std)  # :192  (no *2 in this fork)
    return fb


# ----------------------------------------------------------------------------------
# Whole-item restatement with the reference's RNG consumption order (:139-202)
# ----------------------------------------------------------------------------------
@dataclass
class ItemDraw:
    """Host-drawn random parameters of one __getitem__ call."""
    mixed: bool
    partner: int  # dataset index of the mixup partner, -1 if none
    lam: float  # float64 Beta(10,10) draw, 0.0 if not mixed
    mask: Tuple[int, int, int, int]  # f_start, f_width, t_start, t_width


def item_torch(waves: Sequence[torch.Tensor], index: int, audio_conf: dict, py_random, np_random=np.random,
               use_torchaudio: bool = True, sample_rate: float = SAMPLE_RATE) -> Tuple[torch.Tensor, ItemDraw]:
    """Restatement of ``AudiosetDataset.__getitem__`` (features only, labels live in
    ``psla_b200.dataloaders`` host code) for in-memory waveforms ``waves[i]`` ([1, n_i]).

    RNG order: ``random.random()`` (:147), ``random.randint`` (:153), ``np.random.beta`` (:114),
    then 2 ``torch.rand(1)`` per active mask, freq before time (:182-185); with ``noise``: ``torch.rand(T, F)``,
    ``np.random.rand()``, ``np.random.randint(-10, 10)`` (:198-199)."""
    mixup = audio_conf.get("mixup")
    T = audio_conf.get("target_length")
    F = audio_conf.get("num_mel_bins")
    if py_random.random() < mixup:
        partner = py_random.randint(0, len(waves) - 1)
        lam = float(np_random.beta(10, 10))
        fb = wav2fbank_torch(waves[index], waves[partner], lam, T, F, sample_rate=sample_rate, use_torchaudio=use_torchaudio)
        mixed = True
    else:
        partner, lam, mixed = -1, 0.0, False
        fb = wav2fbank_torch(waves[index], None, 0.0, T, F, sample_rate=sample_rate, use_torchaudio=use_torchaudio)
    mask = draw_mask_offsets(audio_conf.get("freqm"), audio_conf.get("timem"), F, T)
    skip = bool(audio_conf.get("skip_norm")) if audio_conf.get("skip_norm") else False
    out = apply_masks_and_norm_torch(fb, mask, audio_conf.get("mean"), audio_conf.get("std"), skip)
    if audio_conf.get("noise") == True:  # noqa: E712  (audioset_dataset.py:197-199)
        out = noise_roll_torch(out, torch.rand(out.shape[0], out.shape[1]), np_random.rand(), np_random.randint(-10, 10))
    return out, ItemDraw(mixed, partner, lam, mask)


def noise_roll_torch(fbank: torch.Tensor, field: torch.Tensor, scale: float, shift: int) -> torch.Tensor:
    """audioset_dataset.py:198-199 with the three random draws made explicit: `fbank + field * scale / 10`, rolled by
    `shift` rows along time (torch.roll: row t moves to (t + shift) mod T)."""
    return torch.roll(fbank + field * scale / 10, int(shift), 0)


def batch_torch(waves: Sequence[torch.Tensor], src1: Sequence[int], src2: Sequence[int], lam: Sequence[float],
                masks: Sequence[Sequence[int]], target_length: int, norm_mean: float, norm_div: float,
                skip_norm: bool = False, num_mel_bins: int = 128, use_torchaudio: bool = True,
                dtype=torch.float32, sample_rate: float = SAMPLE_RATE) -> torch.Tensor:
    """Batch oracle with explicit parameters (the same arguments the C-ABI takes):
    out[b] = norm(mask(pad(fbank(mix(waves[src1[b]], waves[src2[b]], lam[b])))))."""
    outs = []
    for b in range(len(src1)):
        w1 = waves[src1[b]].to(dtype)
        w2 = waves[src2[b]].to(dtype) if src2[b] >= 0 else None
        if dtype == torch.float32:
            fb = wav2fbank_torch(w1, w2, float(lam[b]), target_length, num_mel_bins, sample_rate=sample_rate, use_torchaudio=use_torchaudio)
        else:
            wave = mix_waveforms_torch(w1, w2, float(lam[b]))
            fb = pad_or_truncate(ref_fbank(wave, num_mel_bins, sample_rate) if use_torchaudio else
                                 kaldi_fbank_torch(wave[0], num_mel_bins, sample_rate), target_length)
        outs.append(apply_masks_and_norm_torch(fb, masks[b], norm_mean, norm_div, skip_norm))
    return torch.stack(outs)
