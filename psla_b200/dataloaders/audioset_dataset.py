"""`AudiosetDataset` with the feature frontend moved to the GPU.

Same constructor, same `audio_conf` keys, same manifest / label-CSV formats, same label logic and the
same RNG consumption order as the reference class (src/dataloaders/audioset_dataset.py:47-205).  The
difference is WHERE the features are computed: a dataloader worker returns the decoded waveform(s) plus
the host-drawn augmentation parameters (`RawItem`); a batch of them collates into one packed buffer
(`RawBatch`); the GPU frontend turns that into the reference's `(audio_input [B, T, F] float32, labels [B, C]
float32)` pair, already on the GPU.

Three ways to consume it, from "no edit" to "fastest":
  1. stock `torch.utils.data.DataLoader(AudiosetDataset(...), batch_size=..., num_workers=..., pin_memory=True)` exactly as
     src/run.py:82-103, src/ensemble/ensemble.py:99-102 and src/ensemble/weight_averaging.py:86-89 build it (no collate_fn:
     `RawItem` registers its own collate with torch's default_collate), and the reference's own loop
     `for i, (audio_input, labels) in enumerate(loader)` (src/traintest.py:81, :245): unpacking a `RawBatch` into two
     names runs the frontend on the current CUDA device (`RawBatch.__iter__`).
  2. `FrontendLoader(loader, frontend)`: the same, with an explicit frontend, the host->device copy of batch k + 1 on a
     side stream while batch k is computed / consumed, and the reference's data-time meters (traintest.py:26-31,87-88).
  3. `FrontendLoader.create(dataset, frontend, batch_size=..., num_workers=...)`: additionally the workers pack their
     batch straight into page-locked shared memory (`PinnedSlotPool`), so the only copy between the decoder and the GPU
     is the DMA; with `loader=pcm16_loader` the workers hand over 16-bit PCM (half the bytes, no float conversion).
Workers never touch CUDA and never pin memory.

File decode stays on the CPU workers (torchaudio.load, :91,95-96) and is pluggable because torchcodec
is not installed in every image: `loader(path) -> (float32 [C, n] tensor in [-1, 1] or int16 PCM [C, n], sample_rate)`.
"""
from __future__ import annotations

import collections
import csv
import functools
import json
import random
import struct
import time
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


def _wav_header(fh):
    """Minimal RIFF/WAVE chunk walk: (format_tag, channels, sample_rate, bits, data_offset, data_bytes)."""
    head = fh.read(12)
    if len(head) < 12 or head[:4] != b"RIFF" or head[8:12] != b"WAVE":
        raise ValueError("not a RIFF/WAVE file")
    fmt = None
    while True:
        ch = fh.read(8)
        if len(ch) < 8:
            raise ValueError("no data chunk")
        size = int.from_bytes(ch[4:8], "little")
        if ch[:4] == b"fmt ":
            body = fh.read(size + (size & 1))
            fmt = struct.unpack("<HHIIHH", body[:16])
        elif ch[:4] == b"data":
            if fmt is None:
                raise ValueError("data chunk before fmt chunk")
            return fmt[0], fmt[1], fmt[2], fmt[5], fh.tell(), size
        else:
            fh.seek(size + (size & 1), 1)


class LazyPcm16:
    """The samples of a mono 16-bit PCM wav file that have NOT been read yet: `pack_waveforms` reads them straight into
    their place in the packed batch buffer (one copy from the page cache into the page-locked slot, none in between)."""
    __slots__ = ("path", "offset", "n")
    dtype = torch.int16

    def __init__(self, path: str, offset: int, n: int):
        self.path, self.offset, self.n = path, int(offset), int(n)

    @property
    def shape(self):
        return (1, self.n)

    def read_into(self, dst: torch.Tensor):
        """dst: contiguous 1-D int16 host tensor of n samples."""
        view = memoryview(dst.numpy()).cast("B")
        with open(self.path, "rb", buffering=0) as fh:
            fh.seek(self.offset)
            got = fh.readinto(view)
        if got != 2 * self.n:
            raise IOError(f"{self.path}: short read ({got} of {2 * self.n} bytes)")

    def materialize(self) -> torch.Tensor:
        x = torch.empty(self.n, dtype=torch.int16)
        self.read_into(x)
        return x.unsqueeze(0)


def pcm16_loader(path: str, lazy: bool = False):
    """16-bit PCM wav file -> (int16 [C, n], sample_rate) with NO float conversion: the GPU kernel applies the decoder's
    value / 32768 itself (psla_frontend_forward_pcm16), so the features equal those of torchaudio.load's floats bit for bit
    while half the bytes cross shared memory and PCIe.  lazy=True returns a `LazyPcm16` for mono files (the samples are
    read when the batch is packed, directly into the packed buffer).  Other encodings fall back to `default_loader`."""
    try:
        with open(path, "rb") as fh:
            tag, ch, sr, bits, off, nbytes = _wav_header(fh)
            if tag != 1 or bits != 16 or ch < 1:
                raise ValueError("not 16-bit PCM")
            n = nbytes // (2 * ch)
            if lazy and ch == 1:
                return LazyPcm16(path, off, n), int(sr)
            data = np.fromfile(fh, dtype="<i2", count=n * ch)
    except (ValueError, struct.error):
        return default_loader(path)
    x = torch.from_numpy(data.astype(np.int16, copy=False)).view(-1, ch).t().contiguous()
    return x, int(sr)


pcm16_lazy_loader = functools.partial(pcm16_loader, lazy=True)


def _in_worker() -> bool:
    return torch.utils.data.get_worker_info() is not None


class PinnedSlotPool:
    """`n_slots` packed-batch buffers in SHARED memory that the main process has page-locked (cudaHostRegister): a
    dataloader worker takes a free slot, packs its batch into it (plain memory writes -- workers never touch CUDA), and the
    main process DMAs straight out of it.  Create it BEFORE the DataLoader starts its workers (they inherit it)."""

    def __init__(self, n_slots: int, slot_samples: int, dtype: torch.dtype = torch.int16, register: Optional[bool] = None,
                 label_floats: int = 0):
        import torch.multiprocessing as mp

        self.n_slots, self.slot_samples, self.dtype = int(n_slots), int(slot_samples), dtype
        self.buf = torch.zeros((self.n_slots, self.slot_samples), dtype=dtype).share_memory_()
        # the batch's label matrix travels in the slot too: a torch tensor sent through the DataLoader's queue costs a
        # round trip to the worker's file-descriptor server (0.5 ms per batch on the main thread, measured)
        self.label_floats = int(label_floats)
        self.labels = torch.zeros((self.n_slots, max(self.label_floats, 1)), dtype=torch.float32).share_memory_()
        self.free = mp.get_context(mp.get_start_method(allow_none=True) or "fork").Queue()
        for i in range(self.n_slots):
            self.free.put(i)
        self.registered = False
        if register is None:
            register = torch.cuda.is_available()
        if register:
            torch.cuda.init()
            err = torch.cuda.cudart().cudaHostRegister(self.buf.data_ptr(), self.buf.numel() * self.buf.element_size(), 0)
            err2 = torch.cuda.cudart().cudaHostRegister(self.labels.data_ptr(), self.labels.numel() * 4, 0)
            self.registered = int(err) == 0 and int(err2) == 0
        self._owner_pid = __import__("os").getpid()

    def view(self, slot: int, n: int) -> torch.Tensor:
        return self.buf[slot, :n]

    def release(self, slot: int):
        self.free.put(int(slot))

    def close(self, n_waiters: int = 64):
        """Unblock workers that wait for a free slot (their batch then travels through ordinary shared memory): call before
        abandoning a loader whose iterator is not exhausted, so that the DataLoader can shut its workers down at once."""
        for _ in range(n_waiters):
            self.free.put(-1)

    def __del__(self):
        try:
            if self.registered and __import__("os").getpid() == self._owner_pid:
                torch.cuda.cudart().cudaHostUnregister(self.buf.data_ptr())
                torch.cuda.cudart().cudaHostUnregister(self.labels.data_ptr())
        except Exception:
            pass


@dataclass
class RawItem:
    wave: torch.Tensor  # [1, n] float32, channel 0 of the primary file
    wave2: Optional[torch.Tensor]  # partner waveform or None
    params: P.ItemParams
    labels: torch.Tensor  # [label_num] float32 (multi-hot, or mixup-soft)
    noise: Optional[torch.Tensor] = None  # [T, F] U[0,1) field of the noise augmentation when it is drawn on the host
    conf: Optional[tuple] = None  # (num_mel_bins, target_length, mean, std, skip_norm, sample_rate) of the dataset that made it


@dataclass
class RawBatch:
    waves: PackedWaves  # host memory: pageable / shared (from a worker), pinned (main process, pin thread), or a pool slot
    src1: np.ndarray
    src2: np.ndarray
    lam: np.ndarray
    masks: np.ndarray
    labels: torch.Tensor  # [B, C]
    noise_scale: Optional[np.ndarray] = None  # [B] float64, set when the noise / roll augmentation is on
    roll: Optional[np.ndarray] = None  # [B] int32
    noise: Optional[torch.Tensor] = None  # [B, T, F] host-drawn fields (pinned), or None: the kernel draws them
    conf: Optional[tuple] = None  # (num_mel_bins, target_length, mean, std, skip_norm, sample_rate)
    slot: int = -1  # PinnedSlotPool slot holding the waveforms (waves.wav is then resolved by the FrontendLoader)
    labels_shape: Optional[tuple] = None  # set when the labels sit in the slot as well (`labels` is then None in transit)

    def __len__(self):
        return len(self.src1)

    def pin_memory(self):
        """What `DataLoader(pin_memory=True)`'s pin thread calls in the MAIN process: page-lock the big buffers there."""
        w = self.waves
        if w.wav is not None and not w.wav.is_pinned():
            self.waves = PackedWaves(w.wav.pin_memory(), w.offsets, w.lengths)
        if self.labels is not None:
            self.labels = self.labels.pin_memory()
        if self.noise is not None:
            self.noise = self.noise.pin_memory()
        return self

    def __iter__(self):
        """`audio_input, labels = batch`: the reference's loops (traintest.py:81, :245) unpack what the loader yields into
        these two names; doing so runs the GPU frontend (a process-wide one per configuration) on the current device."""
        if _in_worker():
            raise RuntimeError("RawBatch is turned into features in the main process (the workers never touch CUDA)")
        fbank, labels = _default_frontend_loader(self.conf).features(self, with_labels=True)
        yield fbank
        yield labels


_DEFAULT_LOADERS = {}


def _default_frontend_loader(conf) -> "FrontendLoader":
    if conf is None:
        raise RuntimeError("this RawBatch carries no frontend configuration; wrap the loader in FrontendLoader(loader, frontend)")
    key = (torch.cuda.current_device(), conf)
    if key not in _DEFAULT_LOADERS:
        nmel, T, mean, std, skip, *rest = conf
        fe = GpuFbankFrontend({"num_mel_bins": nmel, "target_length": T, "mean": mean, "std": std, "skip_norm": skip},
                              sample_rate=rest[0] if rest else 16000)
        _DEFAULT_LOADERS[key] = FrontendLoader(None, fe)
    return _DEFAULT_LOADERS[key]


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
        sr = self.audio_conf.get("sample_rate")  # optional extra key; otherwise the rate of the first file that is loaded
        self.sample_rate = int(sr) if sr is not None else None

    def __len__(self):
        return len(self.data)

    def _load(self, path) -> torch.Tensor:
        # The reference hands the decoder's rate to kaldi.fbank PER FILE (:91,119), which derives the frame geometry from it
        # (kaldi.py:138-140).  A GPU frontend handle has one geometry, so a dataset must be of one rate (every recipe
        # resamples to 16 kHz, egs/*/prep_*.py); the rate travels with the batch and a second rate is an error here.
        wav, sr = self.loader(path)
        if sr is not None:
            if self.sample_rate is None:
                self.sample_rate = int(sr)
            elif int(sr) != self.sample_rate:
                raise ValueError(f"{path}: sample rate {int(sr)} differs from the dataset's {self.sample_rate}; "
                                 "one frontend handles one frame geometry (resample, or use one dataset per rate)")
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
        return RawItem(w1, w2, params, torch.FloatTensor(label_indices), field, self._conf_key())

    def _conf_key(self) -> tuple:
        return (self.melbins, self.audio_conf.get("target_length"), self.norm_mean, self.norm_std, bool(self.skip_norm),
                self.sample_rate if self.sample_rate is not None else 16000)

    def features(self, index: int, frontend: Optional[GpuFbankFrontend] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        """Per-item compatibility with the reference's `dataset[index] -> (fbank [T, F], label_indices)` (:202-205) for code
        that indexes the dataset directly: one clip through the GPU frontend (features stay on the device)."""
        batch = self.collate([self[index]])
        loader = FrontendLoader(None, frontend) if frontend is not None else _default_frontend_loader(batch.conf)
        return loader.features(batch)[0], batch.labels[0]

    # ---- batch assembly ----
    @staticmethod
    def collate(items: Sequence[RawItem], pinned: Optional[bool] = None, pool: Optional[PinnedSlotPool] = None) -> RawBatch:
        """Pack a batch.  In a dataloader WORKER nothing is pinned and CUDA is never touched: the packed buffer goes to the
        main process through shared memory (or is written straight into a `pool` slot); pinning is the main process's job
        (`RawBatch.pin_memory`, called by the DataLoader's pin thread, or the pool's registration).  Called in the main
        process (num_workers=0), `pinned=None` pins when a GPU is there."""
        worker = _in_worker()
        if pinned is None:
            pinned = (not worker) and pool is None and torch.cuda.is_available()
        if worker:
            pinned = False
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
        slot = -1
        dst = None
        if pool is not None:
            total = sum((int(w.shape[-1]) + 3) // 4 * 4 for w in waves)
            all16 = all(w.dtype == torch.int16 for w in waves)
            if total <= pool.slot_samples and ((pool.dtype == torch.int16) == all16):
                slot = pool.free.get()  # back-pressure: blocks until the main process has released one
                if slot >= 0:  # (-1: the pool was closed)
                    dst = pool.view(slot, max(total, 8))
        packed = pack_waveforms(waves, device=None, pinned=pinned, out=dst)
        if slot >= 0:  # the main process has the same mapping: only the slot index travels
            packed = PackedWaves(None, packed.offsets, packed.lengths)
        labels, labels_shape = None, None
        if slot >= 0 and items and len(items) * items[0].labels.numel() <= pool.label_floats:
            labels_shape = (len(items), items[0].labels.numel())
            torch.stack([it.labels for it in items], out=pool.labels[slot, :labels_shape[0] * labels_shape[1]].view(labels_shape))
        else:
            labels = torch.stack([it.labels for it in items])
        batch = RawBatch(packed, np.array(src1, np.int32), np.array(src2, np.int32), np.array(lam, np.float64),
                         np.array(masks, np.int32).reshape(-1, 4), labels, conf=items[0].conf if items else None, slot=slot,
                         labels_shape=labels_shape)
        if any(getattr(it.params, "noise", False) for it in items):
            batch.noise_scale = np.array([it.params.noise_scale for it in items], np.float64)
            batch.roll = np.array([it.params.roll for it in items], np.int32)
            if all(it.noise is not None for it in items):
                field = torch.stack([it.noise for it in items])
                batch.noise = field.pin_memory() if pinned else field
        return batch


def _collate_raw_items(batch, *, collate_fn_map=None):
    return AudiosetDataset.collate(batch)


# a stock DataLoader without collate_fn (the way the reference builds every loader) collates RawItems into a RawBatch
try:
    from torch.utils.data._utils.collate import default_collate_fn_map as _fn_map

    _fn_map[RawItem] = _collate_raw_items
except Exception:  # pragma: no cover  (very old torch: pass collate_fn=AudiosetDataset.collate)
    pass


class _Meter:
    """AverageMeter of the reference (src/utilities/util.py:99-114; used at traintest.py:26-31)."""

    def __init__(self):
        self.val = self.avg = self.sum = 0.0
        self.count = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count


class FrontendLoader:
    """Wraps a DataLoader over `AudiosetDataset` (stock collate, `collate_fn=AudiosetDataset.collate`, or the pooled collate
    of `FrontendLoader.create`); yields `(fbank_cuda, labels_cuda)` -- what traintest.py:81-85 has after its `.to(device)`.

    The packed waveforms of a batch are copied host -> device on a SIDE stream into one of `n_buffers` rotating device
    buffers; the frontend kernel on the caller's current stream waits for that copy by an event (no host synchronisation),
    so the copy of batch k + 1 runs while batch k is transformed and consumed.  A device buffer is reused only after the
    kernel that read it has finished (event), a pool slot is released as soon as its copy has.
    `data_time` / `per_sample_data_time` are the reference's meters (traintest.py:87-88): host seconds from the end of the
    consumer's previous step until the features of the next batch are enqueued."""

    def __init__(self, loader, frontend: GpuFbankFrontend, pool: Optional[PinnedSlotPool] = None, n_buffers: int = 3):
        self.loader = loader
        self.frontend = frontend
        self.pool = pool
        self.n_buffers = int(n_buffers)
        self.data_time, self.per_sample_data_time = _Meter(), _Meter()
        self._copy_stream = None
        self._dev = [None] * self.n_buffers        # uint8 device buffers
        self._dev_free = [None] * self.n_buffers   # event: the kernel that last read the buffer is done
        self._k = 0
        self._pending = collections.deque()        # (copy-done event, pool slot)
        self._label_ring = [None] * self.n_buffers
        self._noise_calls = 0

    @classmethod
    def create(cls, dataset: "AudiosetDataset", frontend: GpuFbankFrontend, batch_size: int = 1, num_workers: int = 0,
               sampler=None, batch_sampler=None, shuffle: bool = False, drop_last: bool = False, max_clip_seconds: float = 10.0,
               pool_slots: Optional[int] = None, pcm16: bool = True, prefetch_factor: Optional[int] = None, **kw) -> "FrontendLoader":
        """DataLoader + pinned shared slot pool in one call (way 3 of the module docstring).  `sampler` / `batch_sampler`
        are passed through (WeightedRandomSampler, psla_b200.shard.ShardedBatchSampler ...).  A slot holds a full batch of
        `max_clip_seconds` clips with every clip mixed; a batch that does not fit travels through ordinary shared memory."""
        bs = batch_size if batch_sampler is None else getattr(batch_sampler, "batch_size", batch_size)
        slot_samples = int(2 * bs * (int(max_clip_seconds * 16000) + 4))
        # The DataLoader hands batches over IN ORDER: up to num_workers * prefetch_factor batches (each holding a slot) can
        # be finished and waiting behind the one that is next in line, and the worker producing THAT one must still find a
        # free slot -- so the pool has to be larger than everything that can be in flight, or the loader deadlocks.
        in_flight = max(1, num_workers) * (prefetch_factor or 2)
        n_slots = max(pool_slots or 0, in_flight + 3)
        pool = PinnedSlotPool(n_slots, slot_samples, torch.int16 if pcm16 else torch.float32, label_floats=bs * int(dataset.label_num))
        args = dict(num_workers=num_workers, collate_fn=functools.partial(AudiosetDataset.collate, pool=pool), **kw)
        if num_workers > 0 and prefetch_factor is not None:
            args["prefetch_factor"] = prefetch_factor
        if batch_sampler is not None:
            dl = torch.utils.data.DataLoader(dataset, batch_sampler=batch_sampler, **args)
        else:
            dl = torch.utils.data.DataLoader(dataset, batch_size=batch_size, sampler=sampler, shuffle=shuffle if sampler is None else False,
                                             drop_last=drop_last, **args)
        return cls(dl, frontend, pool=pool)

    def __len__(self):
        return len(self.loader)

    def close(self):
        """Stop early (iterator not exhausted): wake workers blocked on the slot pool so the DataLoader can shut down."""
        self._reap(block=True)
        if self.pool is not None:
            self.pool.close()

    def host_labels(self, batch: RawBatch) -> torch.Tensor:
        """The batch's [B, C] label matrix on the host: `batch.labels`, or -- for a pooled batch -- the view of its slot
        (valid until the slot is released, i.e. until the batch's host -> device copy has finished)."""
        if batch.labels is not None:
            return batch.labels
        n_lab = batch.labels_shape[0] * batch.labels_shape[1]
        return self.pool.labels[batch.slot, :n_lab].view(batch.labels_shape)

    # ---- one batch ----
    def _reap(self, block: bool = False, keep: int = 0):
        """Release the pool slots whose copies have finished; block=True waits until at most `keep` copies are outstanding."""
        while self._pending and ((block and len(self._pending) > keep) or self._pending[0][0].query()):
            ev, slot = self._pending.popleft()
            ev.synchronize()
            if slot >= 0 and self.pool is not None:
                self.pool.release(slot)

    def _stage(self, batch: RawBatch, with_labels: bool):
        """Host -> device copy of the batch's waveforms (and labels) on the side stream, ordered before whatever the
        caller's current stream does next by an event.  Returns (PackedWaves on the device, labels on the device or None,
        index of the rotating device buffer or None)."""
        dev = self.frontend.device
        wav = batch.waves.wav
        if wav is None:
            if batch.slot < 0 or self.pool is None:
                raise RuntimeError("RawBatch refers to a pool slot but this FrontendLoader has no pool")
            n = int(batch.waves.offsets[-1] + (batch.waves.lengths[-1] + 3) // 4 * 4) if len(batch.waves.lengths) else 8
            wav = self.pool.view(batch.slot, max(n, 8))
        if wav.is_cuda:
            return (PackedWaves(wav, batch.waves.offsets, batch.waves.lengths),
                    batch.labels.to(dev, non_blocking=True) if with_labels else None, None)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        j = self._k % self.n_buffers
        self._k += 1
        nbytes = wav.numel() * wav.element_size()
        cs = self._copy_stream
        main = torch.cuda.current_stream(dev)
        if self._dev[j] is None or self._dev[j].numel() < nbytes:
            # Allocated ON THE COPY STREAM: the caching allocator hands a block that a main-stream tensor has just freed
            # straight back to main-stream allocations, because same-stream reuse is ordered -- but the copy below runs
            # on another stream and would overwrite memory that queued main-stream kernels (the consumer's training
            # step) still use.  (Seen: NaN weights after the third batch of scripts/c5_train.py.)
            if self._dev[j] is not None:
                self._dev[j].record_stream(main)  # its last reader was a kernel on the main stream
            with torch.cuda.stream(cs):
                self._dev[j] = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=dev)
            self._dev_free[j] = None
        dst = self._dev[j][:nbytes].view(wav.dtype)
        labels_dev = None
        lab = self.host_labels(batch)
        if with_labels and not lab.is_pinned():  # small: stage through a page-locked ring so the copy stays asynchronous
            ring = self._label_ring[j]
            if ring is None or ring.numel() < lab.numel() or ring.dtype != lab.dtype:
                ring = self._label_ring[j] = torch.empty(max(lab.numel(), 1), dtype=lab.dtype, pin_memory=True)
            # (the copy that last read this ring entry was reaped n_buffers batches ago at the latest)
            self._reap(block=True, keep=self.n_buffers - 1)
            # (numpy's single-threaded memcpy: a torch copy_ of this size fans out over the intra-op thread pool, whose
            # wake-up costs milliseconds while the dataloader workers keep every core busy -- measured 6 ms vs 0.02 ms)
            ring.numpy()[:lab.numel()] = lab.numpy().reshape(-1)
            lab = ring[:lab.numel()].view(lab.shape)
        if self._dev_free[j] is not None:
            cs.wait_event(self._dev_free[j])
        with torch.cuda.stream(cs):
            dst.copy_(wav, non_blocking=True)
            if with_labels:
                labels_dev = lab.to(dev, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(cs)
        self._pending.append((ev, batch.slot))
        main.wait_event(ev)
        if labels_dev is not None:
            labels_dev.record_stream(main)
        return PackedWaves(dst, batch.waves.offsets, batch.waves.lengths), labels_dev, j

    def features(self, batch: RawBatch, with_labels: bool = False):
        """fbank [B, T, F] on the device (and the labels on the device when `with_labels`), asynchronous on the current stream."""
        self._reap()
        if batch.conf is not None and len(batch.conf) > 5 and int(batch.conf[5]) != self.frontend.sample_rate:
            raise ValueError(f"the batch's files are {int(batch.conf[5])} Hz, the frontend was built for {self.frontend.sample_rate} Hz")
        waves, labels_dev, j = self._stage(batch, with_labels)
        fbank = self.frontend.forward_packed(waves, batch.src1, batch.src2, batch.lam, batch.masks)
        if j is not None:
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(self.frontend.device))
            self._dev_free[j] = done
        if batch.noise_scale is not None:  # audioset_dataset.py:197-199, after the normalisation
            self._noise_calls += 1
            fbank = self.frontend.noise_roll(fbank, batch.noise_scale, batch.roll, batch.noise, seed=self._noise_calls)
        return (fbank, labels_dev) if with_labels else fbank

    def __iter__(self):
        end = time.time()
        it = iter(self.loader)
        try:
            while True:
                # a pool slot goes back to the workers as soon as its copy has finished.  The pool holds 3 slots more than
                # the DataLoader can have in flight: never more than 2 copies are left outstanding while we block in next(),
                # so the worker producing the next batch always finds a free slot
                self._reap(block=self.pool is not None, keep=2)
                try:
                    batch = next(it)
                except StopIteration:
                    break
                fbank, labels = self.features(batch, with_labels=True)
                dt = time.time() - end
                self.data_time.update(dt)
                self.per_sample_data_time.update(dt / max(1, len(batch)))
                yield fbank, labels
                end = time.time()
        finally:
            self._reap(block=True)
