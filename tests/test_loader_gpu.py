"""The worker-fed dataloader path on a B200: real 16-bit wav files -> `torch.utils.data.DataLoader` with
worker processes -> packed host buffer -> side-stream copy -> fused kernel, against golden features of the UNMODIFIED
reference class reading the same files (tests/golden/getitem_files.npz, oracle/gen_golden.py).  `-m gpu`."""
import json
import os
import random

import numpy as np
import pytest
import torch

from oracle import psla_oracle as O
from psla_b200.synth import write_tiny_wav_dataset

pytestmark = pytest.mark.gpu

MEAN, STD, TOL = -4.6476, 4.5699, 1e-3


@pytest.fixture(scope="module")
def files():
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    import tempfile
    with tempfile.TemporaryDirectory(dir=base) as d:  # tmpfs: the decode reads memory, like a warm page cache
        yield write_tiny_wav_dataset(d, copies=2)


@pytest.fixture(scope="module")
def gold(golden_dir):
    return np.load(os.path.join(golden_dir, "getitem_files.npz"))


@pytest.fixture(scope="module")
def D():
    from psla_b200 import build
    build.build()
    import psla_b200.dataloaders as D_
    return D_


def _fill():
    return ((torch.zeros(1) - MEAN) / STD).item()


def _cmp(fb, ref):
    fb, ref = fb.float().cpu(), torch.as_tensor(ref)
    fill = _fill()
    assert torch.equal(fb == fill, ref == fill)  # padding rows / SpecAugment cells: same cells, same constant
    err = ((fb - ref).abs() * STD).max().item()
    assert err <= TOL, err
    return err


def test_reference_loop_over_a_stock_dataloader(D, files, gold):
    """The reference's own construction and loop (run.py:96-98 + traintest.py:245-246), only the package swapped:
    DataLoader(AudiosetDataset(...), batch_size, num_workers, pin_memory=True) and `for i, (audio_input, labels) in ...`."""
    conf = json.loads(str(gold["eval_conf"]))
    val_loader = torch.utils.data.DataLoader(
        D.AudiosetDataset(files[0], label_csv=files[1], audio_conf=conf, verbose=False),
        batch_size=5, shuffle=False, num_workers=4, pin_memory=True)
    device = torch.device("cuda")
    seen = 0
    for i, (audio_input, labels) in enumerate(val_loader):
        audio_input = audio_input.to(device)
        assert audio_input.is_cuda and labels.is_cuda and audio_input.shape[1:] == (256, 128) and audio_input.dtype == torch.float32
        B = audio_input.size(0)
        _cmp(audio_input, gold["eval_feats"][seen:seen + B])
        assert np.array_equal(labels.cpu().numpy(), gold["eval_labels"][seen:seen + B])
        seen += B
    assert seen == 16


def test_pcm16_workers_pinned_pool_and_float_workers_agree(D, files, gold):
    """Way 3 (workers hand over int16 PCM through the page-locked shared pool) == way 2 (float decode, plain collate, pin
    thread) bit for bit, both within 1e-3 of the reference; the data-time meters of traintest.py:87-88 are filled."""
    from psla_b200.frontend import GpuFbankFrontend
    conf = json.loads(str(gold["eval_conf"]))
    fe = GpuFbankFrontend(conf)
    ds16 = D.AudiosetDataset(files[0], conf, label_csv=files[1], loader=D.pcm16_loader, verbose=False)
    fl = D.FrontendLoader.create(ds16, fe, batch_size=6, num_workers=4, max_clip_seconds=2.5)
    assert fl.pool.registered and fl.pool.n_slots == 4 * 2 + 3
    a = []
    for epoch in range(2):
        feats, labs = [], []
        for fb, lab in fl:
            assert fb.is_cuda and lab.is_cuda
            feats.append(fb.clone())
            labs.append(lab.clone())
        a.append(torch.cat(feats))
        assert np.array_equal(torch.cat(labs).cpu().numpy(), gold["eval_labels"])
    assert torch.equal(a[0], a[1])
    _cmp(a[0], gold["eval_feats"])
    assert fl.data_time.count == 6 and fl.per_sample_data_time.avg > 0
    assert fl.pool.free.qsize() == fl.pool.n_slots  # every slot came back
    dsf = D.AudiosetDataset(files[0], conf, label_csv=files[1], verbose=False)  # default loader: float waveforms
    dl = torch.utils.data.DataLoader(dsf, batch_size=6, num_workers=4, collate_fn=D.AudiosetDataset.collate, pin_memory=True)
    b = torch.cat([fb.clone() for fb, _ in D.FrontendLoader(dl, fe)])
    assert torch.equal(a[0], b)
    # per-item compatibility for code that indexes the dataset (reference: dataset[i] -> (fbank, labels), :202-205)
    fb3, lab3 = ds16.features(3, fe)
    assert torch.equal(fb3, a[0][3]) and np.array_equal(lab3.numpy(), gold["eval_labels"][3])


def test_train_configuration_from_files(D, files, gold):
    """Mixup + SpecAugment.  Single process with the generator seeds of oracle/gen_golden.py: the features are the unmodified
    reference's (same draws, same files).  With worker processes every worker has its own RNG streams, so each batch is
    checked against the oracle evaluated with the parameters that batch carries."""
    from psla_b200.frontend import GpuFbankFrontend
    conf = json.loads(str(gold["train_conf"]))
    fe = GpuFbankFrontend(conf)
    ds = D.AudiosetDataset(files[0], conf, label_csv=files[1], loader=D.pcm16_loader, verbose=False)
    random.seed(5); np.random.seed(5); torch.manual_seed(5)
    items = [ds[i] for i in range(len(ds))]  # (a DataLoader iterator would draw its base seed from torch's generator first)
    fl0 = D.FrontendLoader(None, fe)
    got = torch.cat([fl0.features(ds.collate(items[i:i + 4])).clone() for i in range(0, len(items), 4)])
    _cmp(got, gold["train_feats"])
    fl = D.FrontendLoader.create(ds, fe, batch_size=4, num_workers=4, max_clip_seconds=2.5)
    n_mixed = 0
    for batch in fl.loader:
        fb = fl.features(batch)
        wav = fl.pool.view(batch.slot, int(batch.waves.offsets[-1] + batch.waves.lengths[-1]))
        waves = [wav[o:o + n].float().div(32768.0).unsqueeze(0) for o, n in zip(batch.waves.offsets, batch.waves.lengths)]
        ref = O.batch_torch(waves, batch.src1.tolist(), batch.src2.tolist(), batch.lam.tolist(),
                            [tuple(int(v) for v in m) for m in batch.masks], 256, MEAN, STD)
        torch.cuda.synchronize()
        _cmp(fb, ref)
        n_mixed += int((batch.src2 >= 0).sum())
        fl._reap(block=True)
    assert n_mixed > 0


def test_staging_buffers_do_not_reuse_memory_of_queued_consumer_work(D, files, gold):
    """Regression (found by scripts/c5_train.py: NaN weights after the third batch).  The loader copies a batch to the
    device on a SIDE stream.  Its device buffer must not be a block the caching allocator has just taken back from a
    main-stream tensor: kernels queued on the main stream may still use that memory, and nothing orders the side-stream
    copy after them.  Here a long chain of main-stream kernels reads a tensor that is freed while they are still queued;
    the first batches of a fresh loader are staged right behind it; the chain's result must be unaffected."""
    from psla_b200.frontend import GpuFbankFrontend
    conf = json.loads(str(gold["eval_conf"]))
    fe = GpuFbankFrontend(conf)
    ds = D.AudiosetDataset(files[0], conf, label_csv=files[1], verbose=False)  # float waveforms: > 1 MB per batch
    batches = [ds.collate([ds[i] for i in range(16)], pinned=True) for _ in range(3)]
    nbytes = batches[0].waves.wav.numel() * 4
    dev = fe.device
    for rep in range(5):
        fl = D.FrontendLoader(None, fe)  # fresh: its device buffers are allocated at the first three batches
        torch.cuda.synchronize()
        x = torch.ones(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=dev)  # the size the loader will ask for
        acc = torch.zeros(x.numel(), dtype=torch.int32, device=dev)
        big = torch.ones(64 << 20, device=dev)
        for _ in range(60):  # keep the stream busy in front of the reads of x
            big.mul_(1.0000001)
        for _ in range(40):
            acc.add_(x)
        del x  # back to the allocator while the adds are still queued
        feats = [fl.features(b).clone() for b in batches]
        torch.cuda.synchronize()
        assert int(acc.min()) == 40 and int(acc.max()) == 40, rep
        assert torch.equal(feats[0], feats[1]) and torch.equal(feats[0], feats[2])
        _cmp(feats[0], gold["eval_feats"][:16])
