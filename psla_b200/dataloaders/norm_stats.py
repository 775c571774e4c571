"""Dataset normalisation statistics (`norm_mean`, `norm_std` of audio_conf).

The reference points at a `src/get_norm_stats.py` it does not ship (audioset_dataset.py:72-76): run the dataset with
`skip_norm=True` and take mean / std of the un-normalised features.  Here the features come out of the GPU frontend and
the accumulation stays on the GPU (float64 sums; plain torch reductions -- this is bookkeeping, not a kernel).
Padding rows (0.0, audioset_dataset.py:122-128) are part of the tensors the model sees and are counted, as they are
when the statistics are taken over the dataset's outputs.
"""
from __future__ import annotations

from typing import Iterable, Tuple

import torch


class NormStats:
    def __init__(self):
        self.n = 0
        self.s1 = None
        self.s2 = None

    def update(self, fbank: torch.Tensor) -> None:
        x = fbank.double()
        s1, s2 = x.sum(), (x * x).sum()
        self.s1 = s1 if self.s1 is None else self.s1 + s1
        self.s2 = s2 if self.s2 is None else self.s2 + s2
        self.n += fbank.numel()

    def result(self) -> Tuple[float, float]:
        """(mean, std) over every cell seen so far (population std, like `tensor.std(unbiased=False)`)."""
        if self.n == 0:
            raise ValueError("no features seen")
        mean = (self.s1 / self.n).item()
        var = max((self.s2 / self.n).item() - mean * mean, 0.0)
        return mean, var ** 0.5


def compute_norm_stats(feature_batches: Iterable) -> Tuple[float, float]:
    """feature_batches: an iterable of `[B, T, F]` tensors or of `(fbank, labels)` pairs, e.g. a `FrontendLoader` whose
    dataset was built with `skip_norm=True`."""
    st = NormStats()
    for item in feature_batches:
        st.update(item[0] if isinstance(item, (tuple, list)) else item)
    return st.result()
