"""The geometry-generic kernel (psla_b200/csrc/generic_kernel.cuh; SURVEY 8(f4): sample rates other than 16 kHz) through the
C ABI against the oracle and the golden vectors of the unmodified reference.  `-m gpu`.

kaldi.fbank derives the frame geometry from the sample rate (torchaudio/compliance/kaldi.py:138-140): 200 / 80 / 256 at
8 kHz, 551 / 220 / 1024 at 22.05 kHz, 800 / 320 / 1024 at 32 kHz, 1102 / 441 / 2048 at 44.1 kHz, 1200 / 480 / 2048 at
48 kHz.  Tolerance as everywhere: <= 1e-3 abs in log-mel units, fill cells (padding, masks) bit-exact.
"""
import json
import os
import random
import tempfile

import numpy as np
import pytest
import torch

from oracle import psla_oracle as O
from psla_b200.synth import TINY_CLASSES, TINY_LABELS, synth_wave, tiny_dataset_waves

pytestmark = pytest.mark.gpu

MEAN, STD = -4.6476, 4.5699
TOL = 1e-3
RATES = (8000, 22050, 32000, 44100, 48000)


@pytest.fixture(scope="module")
def FE():
    from psla_b200 import build
    build.build()
    from psla_b200.frontend import GpuFbankFrontend
    return GpuFbankFrontend


def _fill(mean=MEAN, div=STD):
    return ((torch.zeros(1) - mean) / div).item()


def _check(out, ref, fill, div=STD, tol=TOL):
    assert out.shape == ref.shape
    a, b = out == fill, ref == fill
    assert int((a ^ b).sum()) <= 1  # a live cell may round to the constant by chance in one of the two
    err = ((out.double() - ref.double()).abs() * div)
    assert err.max().item() <= tol, err.max().item()


def _batch(seed, sr, B=6):
    rng = np.random.default_rng(seed)
    lens = [int(rng.integers(int(0.3 * sr), int(2.5 * sr))) for _ in range(B)]
    waves = [synth_wave(seed * 100 + i, n, ("gauss", "am", "int16")[i % 3]) for i, n in enumerate(lens)]
    waves += [synth_wave(seed * 100 + 50, lens[0] // 2, "gauss"), synth_wave(seed * 100 + 51, lens[2] + 1000, "gauss")]  # shorter / longer partner
    src1 = list(range(B))
    src2 = [B, -1, B + 1, 1, -1, 5][:B]
    lam = [0.37, 0.0, 0.61, 0.5, 0.0, 0.45][:B]
    return waves, src1, src2, lam


@pytest.mark.parametrize("sr", RATES)
def test_other_sample_rates_match_the_oracle(FE, sr):
    shift, size, padded = O.frame_geometry(sr)
    T = 1 + (int(2.0 * sr) - size) // shift  # about two seconds of frames: shorter clips pad, longer ones truncate
    conf = {"num_mel_bins": 128, "target_length": T, "mean": MEAN, "std": STD}
    fe = FE(conf, sample_rate=sr)
    assert (fe.frame_length, fe.frame_shift, fe.fft_size, fe.generic_path) == (size, shift, padded, True)
    waves, src1, src2, lam = _batch(sr % 97, sr)
    masks = [(0, 0, 0, 0), (10, 20, 5, 30), (100, 28, T - 20, 20), (0, 0, 0, 0), (3, 1, 0, 1), (60, 47, 40, 100)]
    ref = O.batch_torch(waves, src1, src2, lam, masks, T, MEAN, STD, sample_rate=sr)
    got = fe(waves, src1, src2, lam, masks).cpu()
    _check(got, ref, _fill())
    assert fe.num_frames(size - 1) == -1 and fe.num_frames(size) == 1 and fe.num_frames(size + shift) == 2
    # 16-bit PCM input: bit-identical to the fp32 path on the decoded floats
    from psla_b200.frontend import to_pcm16
    pcm = [to_pcm16(w) for w in waves]
    dec = [w.to(torch.float32) / 32768.0 for w in pcm]
    assert torch.equal(fe(pcm, src1, src2, lam, masks), fe(dec, src1, src2, lam, masks))
    # skip_norm and an odd number of mel bins
    fe2 = FE({"num_mel_bins": 40, "target_length": T, "skip_norm": True}, sample_rate=sr)
    ref2 = O.batch_torch(waves, src1, src2, lam, [(0, 0, 0, 0)] * 6, T, 0.0, 1.0, skip_norm=True, num_mel_bins=40, sample_rate=sr)
    _check(fe2(waves, src1, src2, lam, None).cpu(), ref2, 0.0, div=1.0)
    fe.close()
    fe2.close()


def test_reference_getitem_at_other_sample_rates(FE, golden_dir):
    """Same seeds -> our AudiosetDataset + the generic kernel reproduce the unmodified reference's __getitem__ output when the
    decoder reports another sample rate (tests/golden/getitem_rates.npz, oracle/gen_golden.py::gen_getitem_rates)."""
    from psla_b200.dataloaders import AudiosetDataset, FrontendLoader
    g = np.load(os.path.join(golden_dir, "getitem_rates.npz"))
    conf = json.loads(str(g["conf"]))
    waves = tiny_dataset_waves()
    store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
    with tempfile.TemporaryDirectory() as tmp:
        jpath, cpath = os.path.join(tmp, "t.json"), os.path.join(tmp, "l.csv")
        json.dump({"data": [{"wav": f"mem://{i}.wav", "labels": TINY_LABELS[i]} for i in range(len(waves))]}, open(jpath, "w"))
        with open(cpath, "w") as f:
            f.write("index,mid,display_name\n" + "".join(f"{i},{m},c{i}\n" for i, m in enumerate(TINY_CLASSES)))
        for sr in g["rates"]:
            sr = int(sr)
            ds = AudiosetDataset(jpath, conf, cpath, loader=lambda p, sr=sr: (store[p].clone(), sr), verbose=False)
            fe = FE(conf, sample_rate=sr)
            random.seed(sr); np.random.seed(sr); torch.manual_seed(sr)
            batch = ds.collate([ds[i] for i in range(len(ds))])
            feats = FrontendLoader(None, fe).features(batch).cpu()
            _check(feats, torch.from_numpy(g[f"feats_{sr}"]), _fill(conf["mean"], conf["std"]), div=conf["std"])
            assert torch.equal(batch.labels, torch.from_numpy(g[f"labels_{sr}"]))
            fe.close()


def test_16k_generic_path_agrees_with_streaming_kernel_and_oracle(FE):
    """Two independent implementations of the recipe (fused streaming kernel vs one-warp-per-frame generic kernel)."""
    T = 320
    conf = {"num_mel_bins": 128, "target_length": T, "mean": MEAN, "std": STD}
    waves, src1, src2, lam = _batch(7, 16000)
    masks = [(0, 0, 0, 0), (10, 20, 5, 30), (100, 28, 300, 20), (0, 0, 0, 0), (3, 1, 0, 1), (60, 47, 40, 100)]
    ref = O.batch_torch(waves, src1, src2, lam, masks, T, MEAN, STD)
    fe = FE(conf, exact_clip_mean=True)
    assert not fe.generic_path
    stream = fe(waves, src1, src2, lam, masks).cpu()
    n0 = fe.launch_count
    fe.set_generic_path(True)
    assert fe.generic_path
    gen = fe(waves, src1, src2, lam, masks).cpu()
    assert fe.launch_count - n0 == 2  # statistics kernel + generic kernel
    fe.set_generic_path(False)
    again = fe(waves, src1, src2, lam, masks).cpu()
    assert torch.equal(again, stream)
    _check(gen, ref, _fill())
    _check(stream, ref, _fill())
    _check(gen, stream, _fill(), tol=2 * TOL)  # each within 1e-3 of the reference (near-floor cells of the 'am' clips): 2e-3 apart at most
    assert torch.equal(gen == _fill(), stream == _fill())
    # more than 128 clips in one call (descriptors through the pinned ring), eval clips
    B = 150
    w = [synth_wave(900 + i, 4000 + 37 * i, "gauss") for i in range(B)]
    fe.set_generic_path(True)
    big = fe(w, list(range(B)), None, None, None).cpu()
    fe.set_generic_path(False)
    _check(big, fe(w, list(range(B)), None, None, None).cpu(), _fill(), tol=2 * TOL)
    fe.close()


def test_generic_path_errors(FE):
    from psla_b200._lib import PslaError, ShortClipError
    sr = 32000
    conf = {"num_mel_bins": 128, "target_length": 64, "mean": MEAN, "std": STD}
    fe = FE(conf, sample_rate=sr)
    with pytest.raises(ShortClipError):  # kaldi.py:142 at this geometry: fewer than 800 samples
        fe([synth_wave(1, 799)], [0], None, None, None)
    assert fe([synth_wave(1, 800)], [0], None, None, None).shape == (1, 64, 128)
    with pytest.raises(PslaError):  # the frequency-major layouts belong to the streaming kernel
        fe.set_output_layout("ft")
        fe([synth_wave(1, 8000)], [0], None, None, None)
    fe.close()
    with pytest.raises(PslaError):  # 25 ms at 200 kHz = 5000 samples: FFT of 8192 points, beyond the generic kernel
        FE(conf, sample_rate=200000)
    with pytest.raises(PslaError):  # 25 ms at 1 kHz = 25 samples: FFT of 32 points
        FE(conf, sample_rate=1000)
