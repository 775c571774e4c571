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

from dataclasses import dataclass
from typing import Callable, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .frontend import SAMPLE_RATE, GpuFbankFrontend, PackedWaves

MIN_SAMPLES = 400  # one 25 ms frame at 16 kHz (other rates: frontend.frame_length): kaldi.fbank asserts on anything shorter (torchaudio/compliance/kaldi.py:142)


def chunk_offsets(n_samples: int, window: int, hop: Optional[int] = None, keep_tail: bool = True,
                  min_samples: int = MIN_SAMPLES) -> Tuple[np.ndarray, np.ndarray]:
    """Start sample and length of every window of a recording of `n_samples`.

    Windows start every `hop` samples (default: `window`, i.e. back to back).  The last, shorter window is kept when it
    holds at least one analysis frame (`keep_tail`), like a short file would be; shorter remainders are dropped."""
    if window < min_samples:
        raise ValueError("window must hold at least one 25 ms frame")
    hop = window if hop is None else int(hop)
    if hop < 1:
        raise ValueError("hop must be positive")
    if n_samples < min_samples:
        return np.zeros(0, dtype=np.int64), np.zeros(0, dtype=np.int64)
    starts = np.arange(0, n_samples, hop, dtype=np.int64)
    lengths = np.minimum(window, n_samples - starts)
    full = lengths == window
    keep = full.copy()
    if keep_tail and (~full).any():  # the first short window covers the rest of the recording; later ones are inside it
        i = int(np.argmax(~full))
        keep[i] = lengths[i] >= min_samples
    return starts[keep], lengths[keep]


@dataclass
class RecordingPrediction:
    """Result of `LongRecordingFrontend.predict`."""
    starts: np.ndarray        # [W] start sample of every window
    per_window: torch.Tensor  # [W, C] model output per window (device tensor)
    clip: Optional[torch.Tensor]  # [C] the recording-level prediction (None when aggregate=None)
    aggregate: Optional[str]

    def top(self, k: int = 5, names: Optional[Sequence[str]] = None) -> List[Tuple[object, float]]:
        """The k strongest classes of the recording: (class index or display name, score)."""
        if self.clip is None:
            raise ValueError("no recording-level prediction (aggregate=None)")
        v, i = torch.topk(self.clip.float().cpu(), min(k, self.clip.numel()))
        return [((names[j] if names is not None else j), float(s)) for s, j in zip(v.tolist(), i.tolist())]

    def timeline(self, class_index: int, sample_rate: int = SAMPLE_RATE) -> List[Tuple[float, float]]:
        """(window start in seconds, score) of one class over the recording."""
        col = self.per_window[:, class_index].float().cpu().tolist()
        return [(float(s) / sample_rate, float(v)) for s, v in zip(self.starts.tolist(), col)]


AGGREGATES = {
    # the reference's models end in a sigmoid per class (Models.py:143-146): per-window outputs are probabilities.
    "mean": lambda y, w: (y * w.unsqueeze(1)).sum(0) / w.sum(),  # frame-weighted average (a short last window counts less)
    "max": lambda y, w: y.max(0).values,                         # "does it occur anywhere in the recording"
}


class LongRecordingFrontend:
    """Chunked eval frontend over one long waveform.

        lr = LongRecordingFrontend(frontend, wave, window_s=10.0)
        for start_sample, feats in lr.batches(256):      # feats: cuda [b, target_length, 128]
            logits = model(feats)
    """

    def __init__(self, frontend: GpuFbankFrontend, wave: torch.Tensor, window_s: float = 10.0, hop_s: Optional[float] = None,
                 sample_rate: Optional[int] = None, keep_tail: bool = True):
        self.frontend = frontend
        sample_rate = int(frontend.sample_rate if sample_rate is None else sample_rate)
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
        self.starts, self.lengths = chunk_offsets(n, window, hop, keep_tail, min_samples=frontend.frame_length)

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
    def predict(self, model: Callable[[torch.Tensor], torch.Tensor], batch: int = 256, aggregate: Optional[str] = "mean",
                overlap_launches: bool = True) -> RecordingPrediction:
        """One-click inference over the recording: features of `batch` windows per launch -> `model` -> per-window outputs
        -> one recording-level prediction.  The model takes [b, target_length, 128] like EffNetAttention.forward
        (src/models/Models.py:104-107) and returns [b, C] (or a tuple whose first element is that, :146).
        aggregate: "mean" (average over windows, weighted by the number of real frames of each window), "max", or None.
        The frontend works one batch AHEAD of the model on a second stream, so the kernel of batch k + 1 runs while the
        model consumes batch k."""
        if aggregate is not None and aggregate not in AGGREGATES:
            raise ValueError(f"aggregate must be one of {sorted(AGGREGATES)} or None")
        dev = self.frontend.device
        outs: List[torch.Tensor] = []
        main = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev) if overlap_launches and len(self) > batch else None
        ranges = [(lo, min(lo + batch, len(self))) for lo in range(0, len(self), batch)]

        def produce(lo, hi):
            if side is None:
                return self.features(lo, hi), None
            side.wait_stream(main)  # (first call: the recording's upload)
            with torch.cuda.stream(side):
                f = self.features(lo, hi)
                ev = torch.cuda.Event()
                ev.record(side)
            return f, ev

        nxt = produce(*ranges[0]) if ranges else None
        for i in range(len(ranges)):
            feats, ev = nxt
            nxt = produce(*ranges[i + 1]) if i + 1 < len(ranges) else None
            if ev is not None:
                main.wait_event(ev)
                feats.record_stream(main)
            y = model(feats)
            outs.append(y[0] if isinstance(y, (tuple, list)) else y)
        per_window = torch.cat(outs, dim=0) if outs else torch.zeros(0, device=dev)
        clip = None
        if aggregate is not None and len(outs):
            fl, fs = self.frontend.frame_length, self.frontend.frame_shift
            frames = np.minimum(np.maximum(1 + (self.lengths - fl) // fs, 0), self.frontend.target_length)
            w = torch.as_tensor(frames, dtype=per_window.dtype, device=per_window.device)
            clip = AGGREGATES[aggregate](per_window, w)
        return RecordingPrediction(self.starts.copy(), per_window, clip, aggregate)
