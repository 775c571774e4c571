"""Host side of the worker-fed dataloader path -- no GPU needed: real 16-bit wav files, stock
`torch.utils.data.DataLoader` with worker processes, the reference's label / RNG order from files, the shared slot pool."""
import json
import os
import random

import numpy as np
import pytest
import torch

from psla_b200.dataloaders import AudiosetDataset, PinnedSlotPool, RawBatch, pcm16_loader
from psla_b200.dataloaders.audioset_dataset import default_loader
from psla_b200.frontend import pack_waveforms, to_pcm16
from psla_b200.synth import tiny_dataset_waves, write_tiny_wav_dataset


@pytest.fixture(scope="module")
def files(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("wavs"))
    return write_tiny_wav_dataset(d, copies=2)


def _conf(golden_dir, which):
    g = np.load(os.path.join(golden_dir, "getitem_files.npz"))
    return g, json.loads(str(g[f"{which}_conf"]))


def test_pcm16_loader_returns_the_file_samples(files):
    jpath, _ = files
    data = json.load(open(jpath))["data"]
    waves = tiny_dataset_waves()
    for i in (0, 3, 5):
        x, sr = pcm16_loader(data[i]["wav"])
        assert sr == 16000 and x.dtype == torch.int16 and x.shape == (1, waves[i].shape[-1])
        assert torch.equal(x, to_pcm16(waves[i]))
        f, _ = default_loader(data[i]["wav"])  # the float decode the reference sees: value / 32768
        assert f.dtype == torch.float32 and torch.equal(f.reshape(-1), x.reshape(-1).float() / 32768.0)


def test_labels_and_rng_order_from_files_match_reference(files, golden_dir):
    """Single process, seeds as in oracle/gen_golden.py: the soft labels of the train configuration are those of the
    unmodified reference reading the same files -- the draws are consumed in its order."""
    g, conf = _conf(golden_dir, "train")
    ds = AudiosetDataset(files[0], conf, label_csv=files[1], loader=pcm16_loader, verbose=False)
    random.seed(5); np.random.seed(5); torch.manual_seed(5)
    for i in range(len(ds)):
        it = ds[i]
        assert it.wave.dtype == torch.int16
        assert np.array_equal(it.labels.numpy(), g["train_labels"][i]), i


@pytest.mark.parametrize("workers", [0, 2])
def test_stock_dataloader_without_collate_fn_yields_raw_batches(files, golden_dir, workers):
    """The reference builds its loaders without a collate_fn (run.py:82-103, ensemble.py:99-102): RawItems collate into a
    RawBatch by themselves; inside worker processes nothing is pinned and CUDA is never touched."""
    g, conf = _conf(golden_dir, "eval")
    ds = AudiosetDataset(files[0], conf, label_csv=files[1], loader=pcm16_loader, verbose=False)
    dl = torch.utils.data.DataLoader(ds, batch_size=5, shuffle=False, num_workers=workers, pin_memory=False)
    seen = 0
    for batch in dl:
        assert isinstance(batch, RawBatch) and batch.slot == -1
        assert batch.waves.wav.dtype == torch.int16 and (workers == 0 and torch.cuda.is_available() or not batch.waves.wav.is_pinned())
        assert batch.conf == (128, 256, -4.6476, 4.5699, False, 16000)  # the files' rate travels with the batch
        B = len(batch)
        assert np.array_equal(batch.labels.numpy(), g["eval_labels"][seen:seen + B])
        assert (batch.src2 == -1).all() and batch.masks.shape == (B, 4) and not batch.masks.any()
        waves = tiny_dataset_waves()
        for b in range(B):
            o, n = batch.waves.offsets[batch.src1[b]], batch.waves.lengths[batch.src1[b]]
            assert o % 4 == 0 and torch.equal(batch.waves.wav[o:o + n], to_pcm16(waves[(seen + b) % 8]).reshape(-1))
        seen += B
    assert seen == len(ds) == 16


def test_one_sample_rate_per_dataset(files, golden_dir):
    """The reference derives the frame geometry per file from the decoder's rate (audioset_dataset.py:91,119); a frontend
    handle has one geometry: the rate travels in the batch's configuration and a second rate in one dataset is an error."""
    g, conf = _conf(golden_dir, "eval")
    waves = tiny_dataset_waves()
    odd = json.load(open(files[0]))["data"][3]["wav"]
    ds = AudiosetDataset(files[0], conf, label_csv=files[1], verbose=False, loader=lambda p: (waves[0].clone(), 44100 if p == odd else 32000))
    assert ds[0].conf[5] == 32000 and ds[1].conf[5] == 32000 and ds.sample_rate == 32000
    with pytest.raises(ValueError, match="sample rate 44100"):
        ds[3]
    ds2 = AudiosetDataset(files[0], dict(conf, sample_rate=16000), label_csv=files[1], verbose=False, loader=lambda p: (waves[0].clone(), 8000))
    with pytest.raises(ValueError, match="sample rate 8000"):
        ds2[0]


def test_pooled_collate_packs_into_shared_slots(files, golden_dir):
    """Workers write their batch into a slot of the shared pool and send only the slot index; the pool applies
    back-pressure (a worker blocks while no slot is free) and the content equals the plain pack."""
    import functools
    _, conf = _conf(golden_dir, "eval")
    ds = AudiosetDataset(files[0], conf, label_csv=files[1], loader=pcm16_loader, verbose=False)
    g, _ = _conf(golden_dir, "eval")
    pool = PinnedSlotPool(2, 2 * 4 * 40004, torch.int16, register=False, label_floats=4 * 10)
    dl = torch.utils.data.DataLoader(ds, batch_size=4, num_workers=2, collate_fn=functools.partial(AudiosetDataset.collate, pool=pool))
    waves = tiny_dataset_waves()
    seen = 0
    for batch in dl:
        assert batch.slot in (0, 1) and batch.waves.wav is None
        assert batch.labels is None and batch.labels_shape == (len(batch), 10)  # the label matrix travels in the slot as well
        assert np.array_equal(pool.labels[batch.slot, :len(batch) * 10].view(len(batch), 10).numpy(), g["eval_labels"][seen:seen + len(batch)])
        ref = pack_waveforms([to_pcm16(waves[(seen + b) % 8]) for b in range(len(batch))])
        n = ref.wav.numel()
        assert np.array_equal(batch.waves.offsets, ref.offsets) and torch.equal(pool.view(batch.slot, n), ref.wav)
        pool.release(batch.slot)  # what FrontendLoader does once the host->device copy has finished
        seen += len(batch)
    assert seen == 16
    # a batch that does not fit a slot (or has another sample type) travels through ordinary shared memory
    small = PinnedSlotPool(1, 64, torch.int16, register=False)
    b = AudiosetDataset.collate([ds[0], ds[1]], pinned=False, pool=small)
    assert b.slot == -1 and b.waves.wav is not None and small.free.get(timeout=1) == 0


def test_batches_mixing_pcm_and_float_waveforms_use_the_decoder_scale():
    """A 16-bit file next to one the float decoder handled (pcm16_loader falls back for other encodings): the packed batch
    is fp32 and the int16 samples are normalised like torchaudio.load does (value / 32768), not copied as raw integers."""
    a = torch.tensor([[100, -200, 32767, -32768, 5]], dtype=torch.int16)
    b = torch.tensor([[0.25, -0.5, 0.125]])
    p = pack_waveforms([a, b])
    assert p.wav.dtype == torch.float32
    assert torch.equal(p.wav[:5], a[0].float() / 32768.0) and torch.equal(p.wav[8:11], b[0])
    assert pack_waveforms([a, a]).wav.dtype == torch.int16
