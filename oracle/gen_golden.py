"""Generate tests/golden/*.npz from the UNMODIFIED reference run in the build container.

TEST INFRASTRUCTURE ONLY.  Needs ``/root/reference`` (read-only mount) and the installed
torchaudio; neither travels to the GPU box, the committed fixtures do.

    python oracle/gen_golden.py            # rewrites tests/golden/

What is pinned (the reference has no tests of its own):
  * fbank_c1.npz      -- ``kaldi.fbank`` as called at audioset_dataset.py:119-120 on the C1 clip
  * tables.npz        -- ``kaldi.get_mel_banks`` nnz + ``torch.hann_window`` as used by the call
  * mask_offsets.npz  -- cells masked by the real Frequency/TimeMasking modules for 200 seeds
  * getitem_*.npz     -- full ``AudiosetDataset.__getitem__`` outputs on a tiny in-memory dataset
  * getitem_rates.npz -- ``__getitem__`` with ``torchaudio.load`` reporting 8 / 22.05 / 32 / 44.1 / 48 kHz (other frame geometries)
  * getitem_files.npz -- the same class reading REAL 16-bit wav files (written by psla_b200.synth.write_tiny_wav_dataset;
                         ``torchaudio.load`` replaced by a wav reader with its int16 normalisation, value / 32768)
"""
from __future__ import annotations

import json
import os
import random
import sys
import tempfile

import numpy as np
import torch
import torchaudio

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, "tests", "golden")
REF_SRC = "/root/reference/src"

sys.path.insert(0, ROOT)
from psla_b200.synth import synth_wave, tiny_dataset_waves, TINY_LABELS, TINY_CLASSES  # noqa: E402


def import_reference_dataset():
    """Import the reference's dataloaders package as is; patch only the file decode."""
    sys.path.insert(0, REF_SRC)
    import dataloaders  # type: ignore  # /root/reference/src/dataloaders/__init__.py:8

    return dataloaders.AudiosetDataset


def write_tiny_dataset(tmp: str, waves):
    store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
    manifest = {"data": [{"wav": f"mem://{i}.wav", "labels": TINY_LABELS[i]} for i in range(len(waves))]}
    jpath = os.path.join(tmp, "tiny.json")
    with open(jpath, "w") as f:
        json.dump(manifest, f)
    cpath = os.path.join(tmp, "labels.csv")
    with open(cpath, "w") as f:
        f.write("index,mid,display_name\n")
        for i, mid in enumerate(TINY_CLASSES):
            f.write(f"{i},{mid},class{i}\n")
    # torchcodec is absent: torchaudio.load cannot decode here.  The decode is outside the
    # hot path, so serve the in-memory waveforms instead.
    torchaudio.load = lambda path: (store[path].clone(), 16000)
    return jpath, cpath


def gen_fbank_c1():
    x = synth_wave(0, 160000)
    kw = dict(htk_compat=True, sample_frequency=16000, use_energy=False, window_type="hanning",
              num_mel_bins=128, dither=0.0, frame_shift=10)
    fb32 = torchaudio.compliance.kaldi.fbank(x, **kw)
    fb64 = torchaudio.compliance.kaldi.fbank(x.double(), **kw)
    rows = np.r_[0:16, 491:507, 982:998]
    np.savez_compressed(
        os.path.join(GOLD, "fbank_c1.npz"),
        rows=rows, fb32_rows=fb32[rows].numpy(), fb64_rows=fb64[rows].numpy(),
        row_sum64=fb32.double().sum(1).numpy(), col_sum64=fb32.double().sum(0).numpy(),
        shape=np.array(fb32.shape), wave_head=x[0, :16].numpy(),
        max_abs_32_vs_64=np.array((fb32.double() - fb64).abs().max().item()))
    print("fbank_c1", tuple(fb32.shape), "max|32-64| =", (fb32.double() - fb64).abs().max().item())


def gen_tables():
    from torchaudio.compliance import kaldi

    mel, _ = kaldi.get_mel_banks(128, 512, 16000.0, 20.0, 0.0, 100.0, -500.0, 1.0)
    mel = mel.numpy()
    r, c = np.nonzero(mel)
    win = torch.hann_window(400, periodic=False).numpy()
    np.savez_compressed(os.path.join(GOLD, "tables.npz"), mel_rows=r.astype(np.int16), mel_cols=c.astype(np.int16),
                        mel_vals=mel[r, c], mel_shape=np.array(mel.shape), hann400=win)
    print("tables: nnz", len(r), "empty rows", [int(i) for i in range(128) if not (r == i).any()])


def gen_mask_offsets():
    T, F = 1056, 128
    out = np.zeros((200, 4), dtype=np.int32)
    for s in range(200):
        torch.manual_seed(s)
        spec = torch.ones(1, F, T)
        spec = torchaudio.transforms.FrequencyMasking(48)(spec)  # audioset_dataset.py:177,183
        fcols = (spec[0, :, 0] == 0).nonzero().flatten()
        spec2 = torchaudio.transforms.TimeMasking(192)(torch.ones(1, F, T))  # :178,185
        tcols = (spec2[0, 0, :] == 0).nonzero().flatten()
        f0, fw = (int(fcols[0]), len(fcols)) if len(fcols) else (0, 0)
        t0, tw = (int(tcols[0]), len(tcols)) if len(tcols) else (0, 0)
        assert fw == 0 or int(fcols[-1]) == f0 + fw - 1
        assert tw == 0 or int(tcols[-1]) == t0 + tw - 1
        out[s] = (f0, fw, t0, tw)
    np.savez_compressed(os.path.join(GOLD, "mask_offsets.npz"), offsets=out, T=np.array(T), F=np.array(F),
                        freqm=np.array(48), timem=np.array(192))
    print("mask_offsets: width ranges", out[:, 1].min(), out[:, 1].max(), out[:, 3].min(), out[:, 3].max())


def gen_getitem(AudiosetDataset, only=None):
    waves = tiny_dataset_waves()
    with tempfile.TemporaryDirectory() as tmp:
        jpath, cpath = write_tiny_dataset(tmp, waves)
        confs = {
            # the noise / roll augmentation (:197-199), off in every recipe of the reference but part of its dataset API
            "noise": {"num_mel_bins": 128, "target_length": 96, "freqm": 24, "timem": 40, "mixup": 0.5,
                      "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": True},
            "train": {"num_mel_bins": 128, "target_length": 96, "freqm": 24, "timem": 40, "mixup": 0.5,
                      "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False},
            "eval": {"num_mel_bins": 128, "target_length": 96, "freqm": 0, "timem": 0, "mixup": 0,
                     "dataset": "audioset", "mode": "evaluation", "mean": -4.6476, "std": 4.5699, "noise": False},
            "skipnorm": {"num_mel_bins": 128, "target_length": 64, "freqm": 0, "timem": 16, "mixup": 1.0,
                         "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False,
                         "skip_norm": True},
        }
        for name, conf in confs.items():
            if only is not None and name not in only:
                continue
            ds = AudiosetDataset(jpath, audio_conf=conf, label_csv=cpath)
            feats, labels = [], []
            for seed in (0, 1):
                random.seed(seed)
                np.random.seed(seed)
                torch.manual_seed(seed)
                for i in range(len(ds)):
                    fb, lab = ds[i]
                    feats.append(fb.numpy())
                    labels.append(lab.numpy())
            np.savez_compressed(os.path.join(GOLD, f"getitem_{name}.npz"), feats=np.stack(feats),
                                labels=np.stack(labels), conf=json.dumps(conf), seeds=np.array([0, 1]))
            print("getitem", name, np.stack(feats).shape)


RATES = (8000, 22050, 32000, 44100, 48000)


def gen_getitem_rates(AudiosetDataset):
    """The unmodified reference class on waveforms at OTHER sample rates: `torchaudio.load` hands (wave, sr) back and the
    reference passes `sample_frequency=sr` on to kaldi.fbank (audioset_dataset.py:91,119), which derives the frame
    geometry from it (kaldi.py:138-140: 200/80/256 at 8 kHz ... 1200/480/2048 at 48 kHz).  Train conf, seeded."""
    waves = tiny_dataset_waves()
    conf = {"num_mel_bins": 128, "target_length": 48, "freqm": 24, "timem": 12, "mixup": 0.5,
            "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        jpath, cpath = write_tiny_dataset(tmp, waves)
        store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
        for sr in RATES:
            torchaudio.load = lambda path, sr=sr: (store[path].clone(), sr)
            ds = AudiosetDataset(jpath, audio_conf=conf, label_csv=cpath)
            feats, labels = [], []
            random.seed(sr)
            np.random.seed(sr)
            torch.manual_seed(sr)
            for i in range(len(ds)):
                fb, lab = ds[i]
                feats.append(fb.numpy())
                labels.append(lab.numpy())
            out[f"feats_{sr}"] = np.stack(feats)
            out[f"labels_{sr}"] = np.stack(labels)
            print("getitem_rates", sr, out[f"feats_{sr}"].shape)
    np.savez_compressed(os.path.join(GOLD, "getitem_rates.npz"), conf=json.dumps(conf), rates=np.array(RATES), **out)


def gen_getitem_files(AudiosetDataset):
    """The unmodified reference class over real wav files: eval conf (deterministic) and train conf (seeded, single process)."""
    import wave

    from psla_b200.synth import write_tiny_wav_dataset

    def load_wav(path):  # what torchaudio.load(path) returns for a mono 16-bit PCM file (normalize=True)
        with wave.open(path, "rb") as f:
            assert f.getsampwidth() == 2 and f.getnchannels() == 1
            sr = f.getframerate()
            data = np.frombuffer(f.readframes(f.getnframes()), dtype="<i2")
        return torch.from_numpy(data.astype(np.float32) / 32768.0).unsqueeze(0), sr

    with tempfile.TemporaryDirectory() as tmp:
        jpath, cpath = write_tiny_wav_dataset(tmp, copies=2)
        torchaudio.load = load_wav
        out = {}
        confs = {
            "eval": {"num_mel_bins": 128, "target_length": 256, "freqm": 0, "timem": 0, "mixup": 0,
                     "dataset": "audioset", "mode": "evaluation", "mean": -4.6476, "std": 4.5699, "noise": False},
            "train": {"num_mel_bins": 128, "target_length": 256, "freqm": 24, "timem": 40, "mixup": 0.5,
                      "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False},
        }
        for name, conf in confs.items():
            ds = AudiosetDataset(jpath, audio_conf=conf, label_csv=cpath)
            random.seed(5)
            np.random.seed(5)
            torch.manual_seed(5)
            feats, labels = [], []
            for i in range(len(ds)):
                fb, lab = ds[i]
                feats.append(fb.numpy())
                labels.append(lab.numpy())
            out[f"{name}_feats"], out[f"{name}_labels"], out[f"{name}_conf"] = np.stack(feats), np.stack(labels), json.dumps(conf)
        np.savez_compressed(os.path.join(GOLD, "getitem_files.npz"), seed=np.array(5), copies=np.array(2), **out)
        print("getitem_files", out["eval_feats"].shape, out["train_feats"].shape)


def main():
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(1)
    if len(sys.argv) > 1:  # e.g. `python oracle/gen_golden.py noise`: (re)generate only the named __getitem__ fixtures
        if "files" in sys.argv[1:]:
            gen_getitem_files(import_reference_dataset())
        if "rates" in sys.argv[1:]:
            gen_getitem_rates(import_reference_dataset())
        rest = set(sys.argv[1:]) - {"files", "rates"}
        if rest:
            gen_getitem(import_reference_dataset(), only=rest)
        return
    gen_fbank_c1()
    gen_tables()
    gen_mask_offsets()
    gen_getitem(import_reference_dataset())
    gen_getitem_files(import_reference_dataset())
    gen_getitem_rates(import_reference_dataset())
    meta = {"torch": torch.__version__, "torchaudio": torchaudio.__version__, "numpy": np.__version__,
            "reference": "/root/reference (haoheliu/psla), src/dataloaders/audioset_dataset.py unmodified",
            "pinned_in_reference_requirements": "torchaudio==0.6.0 torch==1.6.0 (not installable here)"}
    with open(os.path.join(GOLD, "META.json"), "w") as f:
        json.dump(meta, f, indent=1)


if __name__ == "__main__":
    main()
