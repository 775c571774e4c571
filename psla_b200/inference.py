"""Frontend of the "one-click inference" path for long recordings (README.md:267-268 of the reference lists it as TBD;
BASELINE.json config 4): an hour-long 16 kHz recording is cut into 10 s windows, every window is an independent eval
clip -- its own DC removal (src/dataloaders/audioset_dataset.py:92), no mixup, no masks, padded to target_length
(:122-132), normalised (:191-192) -- exactly what `AudiosetDataset.__getitem__` would return for a file holding that
window in eval mode.

The recording is uploaded ONCE; windows are (offset, length) descriptors into that one device buffer (no per-window
copies, overlapping hops allowed), processed `batch` windows per kernel launch.  The model that consumes the features
(src/models/Models.py:104-107) stays stock PyTorch and is passed in as a callable.
"""
from __future__ import annotations

from typing import Callable, Iterator, List, Optional, Tuple

import numpy as np
import torch

from .frontend import SAMPLE_RATE, GpuFbankFrontend, PackedWaves

MIN_SAMPLES = 400  # one 25 ms frame: kaldi.fbank asserts on anything shorter (torchaudio/compliance/kaldi.py:142)


def chunk_offsets(n_samples: int, window: int, hop: Optional[int] = None, keep_tail: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """Start sample and length of every window of a recording of `n_samples`.

    Windows start every `hop` samples (default: `window`, i.e. back to back).  The last, shorter window is kept when it
    holds at least one analysis frame (`keep_tail`), like a short file would be; shorter remainders are dropped."""
    if window < MIN_SAMPLES:
        raise ValueError("window must hold at least one 25 ms frame")
    hop = window if hop is None else int(hop)
    if hop < 1:
        raise ValueError("hop must be positive")
    if n_samples < MIN_SAMPLES:
        return np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64)
    starts = np.arange(0, n_samples, hop, dtype=np.int64)
    lengths = np.minimum(window, n_samples - starts)
    full = lengths == window
    keep = full.copy()
    if keep_tail and (~full).any():  # the first short window covers the rest of the recording; later ones are inside it
        i = int(np.argmax(~full))
        keep[i] = lengths[i] >= MIN_SAMPLES
    return starts[keep], lengths[keep]


class LongRecordingFrontend:
    """Chunked eval frontend over one long waveform.

        lr = LongRecordingFrontend(frontend, wave, window_s=10.0)
        for start_sample, feats in lr.batches(256):      # feats: cuda [b, target_length, 128]
            logits = model(feats)
    """

    def __init__(self, frontend: GpuFbankFrontend, wave: torch.Tensor, window_s: float = 10.0, hop_s: Optional[float] = None,
                 sample_rate: int = SAMPLE_RATE, keep_tail: bool = True):
        self.frontend = frontend
        w = wave.detach()
        w = w[0] if w.dim() == 2 else w  # channel 0, as kaldi.fbank selects it (kaldi.py:136)
        if w.dtype not in (torch.float32, torch.int16):
            w = w.to(torch.float32)
        n = int(w.shape[0])
        pad = (-n) % 8  # vector loads may touch the rounded-up end of the buffer
        if pad or not w.is_cuda:
            buf = torch.zeros(n + pad + 8, dtype=w.dtype, device=frontend.device)
            buf[:n] = w.to(frontend.device, non_blocking=True)
            w = buf
        self.wave = w
        self.n_samples = n
        window = int(round(window_s * sample_rate))
        hop = None if hop_s is None else int(round(hop_s * sample_rate))
        self.starts, self.lengths = chunk_offsets(n, window, hop, keep_tail)

    def __len__(self) -> int:
        return len(self.starts)

    def features(self, lo: int, hi: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Features of windows [lo, hi) in one launch."""
        packed = PackedWaves(self.wave, self.starts[lo:hi], self.lengths[lo:hi])
        return self.frontend.forward_packed(packed, np.arange(hi - lo, dtype=np.int32), None, None, None, out=out)

    def batches(self, batch: int) -> Iterator[Tuple[np.ndarray, torch.Tensor]]:
        for lo in range(0, len(self), batch):
            hi = min(lo + batch, len(self))
            yield self.starts[lo:hi], self.features(lo, hi)

    @torch.no_grad()
    def predict(self, model: Callable[[torch.Tensor], torch.Tensor], batch: int = 256) -> Tuple[np.ndarray, torch.Tensor]:
        """Window start samples and per-window model outputs (the model takes [b, target_length, 128] like
        EffNetAttention.forward, src/models/Models.py:104-107); aggregation over the recording is the caller's choice."""
        outs: List[torch.Tensor] = []
        for _, feats in self.batches(batch):
            y = model(feats)
            outs.append(y[0] if isinstance(y, (tuple, list)) else y)
        per_window = torch.cat(outs, dim=0) if outs else torch.zeros(0)
        return self.starts.copy(), per_window
