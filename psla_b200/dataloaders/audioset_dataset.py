"""`AudiosetDataset` with the feature frontend moved to the GPU.

Same constructor, same `audio_conf` keys, same manifest / label-CSV formats, same label logic and the
same RNG consumption order as the reference class (src/dataloaders/audioset_dataset.py:47-205).  The
difference is WHERE the features are computed: a dataloader worker returns the decoded waveform(s) plus
the host-drawn augmentation parameters (`RawItem`); `AudiosetDataset.collate` packs a batch into pinned
memory (`RawBatch`); `FrontendLoader` runs `GpuFbankFrontend` on it and yields the reference's
`(audio_input [B, T, F] float32, labels [B, C] float32)` pair, already on the GPU, so it replaces
`for i, (audio_input, labels) in enumerate(train_loader)` at src/traintest.py:81-85 as is.

File decode stays on the CPU workers (torchaudio.load, :91,95-96) and is pluggable because torchcodec
is not installed in every image: `loader(path) -> (float32 [C, n] tensor in [-1, 1], sample_rate)`.
"""
from __future__ import annotations

import csv
import json
import random
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset

from .. import params as P
from ..frontend import GpuFbankFrontend, PackedWaves, pack_waveforms


def make_index_dict(label_csv):  # reference :12-20
    lookup = {}
    with open(label_csv, "r") as f:
        for row in csv.DictReader(f):
            lookup[row["mid"]] = row["index"]
    return lookup


def make_name_dict(label_csv):  # reference :22-30
    lookup = {}
    with open(label_csv, "r") as f:
        for row in csv.DictReader(f):
            lookup[row["index"]] = row["display_name"]
    return lookup


def lookup_list(index_list, label_csv):  # reference :32-37
    table = make_name_dict(label_csv)
    return [table[item] for item in index_list]


def default_loader(path: str) -> Tuple[torch.Tensor, int]:
    """torchaudio.load when it works, else a scipy PCM-wav reader normalised the same way (/32768)."""
    try:
        import torchaudio

        return torchaudio.load(path)
    except Exception:
        from scipy.io import wavfile

        sr, data = wavfile.read(path)
        x = torch.from_numpy(np.asarray(data))
        if x.dtype == torch.int16:
            x = x.to(torch.float32) / 32768.0
        elif x.dtype == torch.int32:
            x = x.to(torch.float32) / 2147483648.0
        else:
            x = x.to(torch.float32)
        x = x.unsqueeze(0) if x.dim() == 1 else x.t().contiguous()
        return x, int(sr)


@dataclass
class RawItem:
    wave: torch.Tensor  # [1, n] float32, channel 0 of the primary file
    wave2: Optional[torch.Tensor]  # partner waveform or None
    params: P.ItemParams
    labels: torch.Tensor  # [label_num] float32 (multi-hot, or mixup-soft)
    noise: Optional[torch.Tensor] = None  # [T, F] U[0,1) field of the noise augmentation when it is drawn on the host


@dataclass
class RawBatch:
    waves: PackedWaves  # pinned host memory
    src1: np.ndarray
    src2: np.ndarray
    lam: np.ndarray
    masks: np.ndarray
    labels: torch.Tensor  # [B, C]
    noise_scale: Optional[np.ndarray] = None  # [B] float64, set when the noise / roll augmentation is on
    roll: Optional[np.ndarray] = None  # [B] int32
    noise: Optional[torch.Tensor] = None  # [B, T, F] host-drawn fields (pinned), or None: the kernel draws them

    def __len__(self):
        return len(self.src1)


class AudiosetDataset(Dataset):
    def __init__(self, dataset_json_file, audio_conf, label_csv=None, loader: Optional[Callable] = None, verbose=True):
        self.datapath = dataset_json_file
        with open(dataset_json_file, "r") as fp:
            self.data = json.load(fp)["data"]
        self.audio_conf = audio_conf
        say = print if verbose else (lambda *a, **k: None)
        say("---------------the {:s} dataloader---------------".format(str(self.audio_conf.get("mode"))))
        self.melbins = self.audio_conf.get("num_mel_bins")
        self.freqm = self.audio_conf.get("freqm")
        self.timem = self.audio_conf.get("timem")
        say("now using following mask: {:d} freq, {:d} time".format(self.freqm, self.timem))
        self.mixup = self.audio_conf.get("mixup")
        say("now using mix-up with rate {:f}".format(self.mixup))
        self.dataset = self.audio_conf.get("dataset")
        self.norm_mean = self.audio_conf.get("mean")
        self.norm_std = self.audio_conf.get("std")
        self.skip_norm = self.audio_conf.get("skip_norm") if self.audio_conf.get("skip_norm") else False
        self.noise = self.audio_conf.get("noise")
        # noise / roll augmentation (:197-199; off in every recipe, run.py:77,80).  The three draws always happen on the
        # host in the reference's order, so the RNG streams stay aligned with the reference; `noise_field` (an extra
        # key, not in the reference) says where the U[0,1) field comes from: "host" ships the drawn torch.rand(T, F)
        # to the GPU (bit-exact), "device" drops it and lets the kernel draw its own (same distribution, no 540 KB per clip).
        self.noise_field = self.audio_conf.get("noise_field", "host")
        if self.noise_field not in ("host", "device"):
            raise ValueError("audio_conf['noise_field'] must be 'host' or 'device'")
        self.index_dict = make_index_dict(label_csv)
        self.label_num = len(self.index_dict)
        say("number of classes is {:d}".format(self.label_num))
        self.loader = loader or default_loader

    def __len__(self):
        return len(self.data)

    def _load(self, path) -> torch.Tensor:
        wav, _sr = self.loader(path)
        return wav

    def __getitem__(self, index) -> RawItem:
        """Same control flow and RNG order as the reference __getitem__ (:139-202); returns raw material."""
        datum = self.data[index]
        mixed, partner = P.draw_mixup(self.mixup, len(self.data), random)
        label_indices = np.zeros(self.label_num)
        if mixed:
            mix_datum = self.data[partner]
            w1, w2 = self._load(datum["wav"]), self._load(mix_datum["wav"])
            lam = P.draw_lambda(np.random)  # drawn inside _wav2fbank, after both loads (:114)
            for label_str in datum["labels"].split(","):
                label_indices[int(self.index_dict[label_str])] += lam
            for label_str in mix_datum["labels"].split(","):
                label_indices[int(self.index_dict[label_str])] += (1.0 - lam)
        else:
            w1, w2, lam = self._load(datum["wav"]), None, 0.0
            for label_str in datum["labels"].split(","):
                label_indices[int(self.index_dict[label_str])] = 1.0
        mask = P.draw_masks(self.freqm, self.timem, self.melbins, self.audio_conf.get("target_length"))
        params = P.ItemParams(mixed, partner, lam, mask)
        field = None
        if self.noise == True:  # noqa: E712  (the reference compares with ==, :197)
            field, params.noise_scale, params.roll = P.draw_noise(self.audio_conf.get("target_length"), self.melbins, np.random)
            if self.noise_field == "device":
                field = None
            params.noise = True
        return RawItem(w1, w2, params, torch.FloatTensor(label_indices), field)

    # ---- batch assembly ----
    @staticmethod
    def collate(items: Sequence[RawItem], pinned: Optional[bool] = None) -> RawBatch:
        if pinned is None:
            pinned = torch.cuda.is_available()
        waves: List[torch.Tensor] = []
        src1, src2, lam, masks = [], [], [], []
        for it in items:
            # the reference takes the mean over ALL channels but fbank reads channel 0 only (:92 vs kaldi.py:136);
            # mono files (every PSLA recipe resamples to 16 kHz mono) make the two coincide
            src1.append(len(waves))
            waves.append(it.wave)
            if it.wave2 is not None:
                src2.append(len(waves))
                waves.append(it.wave2)
            else:
                src2.append(-1)
            lam.append(it.params.lam)
            masks.append(it.params.mask)
        packed = pack_waveforms(waves, device=None, pinned=pinned)
        batch = RawBatch(packed, np.array(src1, np.int32), np.array(src2, np.int32), np.array(lam, np.float64),
                         np.array(masks, np.int32).reshape(-1, 4), torch.stack([it.labels for it in items]))
        if any(getattr(it.params, "noise", False) for it in items):
            batch.noise_scale = np.array([it.params.noise_scale for it in items], np.float64)
            batch.roll = np.array([it.params.roll for it in items], np.int32)
            if all(it.noise is not None for it in items):
                field = torch.stack([it.noise for it in items])
                batch.noise = field.pin_memory() if pinned else field
        return batch


class FrontendLoader:
    """Wraps a DataLoader built with `collate_fn=AudiosetDataset.collate`; yields `(fbank_cuda, labels_cuda)`."""

    def __init__(self, loader, frontend: GpuFbankFrontend):
        self.loader = loader
        self.frontend = frontend

    def __len__(self):
        return len(self.loader)

    def features(self, batch: RawBatch) -> torch.Tensor:
        dev = self.frontend.device
        waves = PackedWaves(batch.waves.wav.to(dev, non_blocking=True), batch.waves.offsets, batch.waves.lengths)
        fbank = self.frontend.forward_packed(waves, batch.src1, batch.src2, batch.lam, batch.masks)
        if batch.noise_scale is not None:  # audioset_dataset.py:197-199, after the normalisation
            self._noise_calls = getattr(self, "_noise_calls", 0) + 1
            fbank = self.frontend.noise_roll(fbank, batch.noise_scale, batch.roll, batch.noise, seed=self._noise_calls)
        return fbank

    def __iter__(self):
        for batch in self.loader:
            yield self.features(batch), batch.labels.to(self.frontend.device, non_blocking=True)
