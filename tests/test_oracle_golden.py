"""Pin the oracle (oracle/psla_oracle.py) against outputs of the UNMODIFIED reference
(tests/golden/*.npz, written by oracle/gen_golden.py in the build container)."""
import json
import os
import random

import numpy as np
import pytest
import torch

from oracle import psla_oracle as O
from psla_b200.synth import synth_wave, tiny_dataset_waves

# The fixtures were produced with MKL's FFT on one host; pocketfft / another MKL build may
# differ by fp32 ulps in the spectrum, which the log turns into <=1e-4 on broadband input.
HOST_DRIFT = 1e-4


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_wave_generator_is_reproducible(golden_dir):
    g = _load(golden_dir, "fbank_c1.npz")
    x = synth_wave(0, 160000)
    assert np.array_equal(x[0, :16].numpy(), g["wave_head"])


def test_tables_match_reference(golden_dir):
    g = _load(golden_dir, "tables.npz")
    mel = O.mel_banks_torch().numpy()
    assert mel.shape == tuple(g["mel_shape"])
    r, c = np.nonzero(mel)
    assert np.array_equal(r, g["mel_rows"]) and np.array_equal(c, g["mel_cols"])
    # same torch ops, same host family: allow 1 ulp of the largest intermediate (mel ~ 2840)
    assert np.abs(mel[r, c] - g["mel_vals"]).max() <= 2e-5
    assert len(r) == 504 and not (r == 3).any()  # row 3 is empty
    assert np.abs(O.hann_window_np(400, np.float32) - g["hann400"]).max() <= 1.2e-7
    # naive fp64 weights differ from the fp32-built ones by ~1e-5 (why the table is built in fp32 order)
    d = np.abs(O.mel_banks_np(dtype=np.float64) - mel.astype(np.float64)).max()
    assert 1e-7 < d < 1e-4


@pytest.mark.parametrize("flavour", ["torchaudio", "torch", "np32", "np64"])
def test_fbank_c1(golden_dir, flavour):
    g = _load(golden_dir, "fbank_c1.npz")
    x = synth_wave(0, 160000)
    if flavour == "torchaudio":
        fb = O.ref_fbank(x).double().numpy()
    elif flavour == "torch":
        fb = O.kaldi_fbank_torch(x[0]).double().numpy()
    elif flavour == "np32":
        fb = O.kaldi_fbank_np(x[0].numpy(), dtype=np.float32).astype(np.float64)
    else:
        fb = O.kaldi_fbank_np(x[0].numpy(), dtype=np.float64)
    assert fb.shape == tuple(g["shape"])
    rows = g["rows"]
    if flavour == "np64":
        assert np.abs(fb[rows] - g["fb64_rows"]).max() <= 1e-9
        # the fp32 reference is itself ~4e-4 from fp64 truth on this clip
        assert np.abs(fb[rows] - g["fb32_rows"]).max() <= 1e-3
    else:
        tol = HOST_DRIFT if flavour != "np32" else 1e-3  # np32: pocketfft fp32, another rounding pattern
        assert np.abs(fb[rows] - g["fb32_rows"]).max() <= tol
        assert np.abs(fb.sum(1) - g["row_sum64"]).max() <= 128 * tol
        assert np.abs(fb.sum(0) - g["col_sum64"]).max() <= 998 * tol


def test_mask_offsets_bit_exact(golden_dir):
    g = _load(golden_dir, "mask_offsets.npz")
    T, F = int(g["T"]), int(g["F"])
    for s in range(200):
        torch.manual_seed(s)
        f0, fw, t0, tw = O.draw_mask_offsets(int(g["freqm"]), int(g["timem"]), F, T)
        # an empty mask leaves no trace in the reference output, so its start is not observable
        got = (f0 if fw else 0, fw, t0 if tw else 0, tw)
        assert got == tuple(int(v) for v in g["offsets"][s]), s


@pytest.mark.parametrize("name", ["train", "eval", "skipnorm"])
@pytest.mark.parametrize("use_torchaudio", [True, False])
def test_getitem_restatement(golden_dir, name, use_torchaudio):
    g = _load(golden_dir, f"getitem_{name}.npz")
    conf = json.loads(str(g["conf"]))
    waves = tiny_dataset_waves()
    k = 0
    for seed in g["seeds"]:
        random.seed(int(seed))
        np.random.seed(int(seed))
        torch.manual_seed(int(seed))
        for i in range(len(waves)):
            out, draw = O.item_torch(waves, i, conf, random, np.random, use_torchaudio=use_torchaudio)
            ref = g["feats"][k]
            assert out.shape == ref.shape
            # masked / padded cells must be the identical fp32 constant, same cells
            fill = np.float32(0.0) if conf.get("skip_norm") else ((torch.zeros(1) - conf["mean"]) / conf["std"]).numpy()[0]
            assert np.array_equal(ref == fill, out.numpy() == fill), (name, k)
            assert np.abs(out.numpy() - ref).max() <= HOST_DRIFT, (name, k)
            k += 1
    assert k == len(g["feats"])


@pytest.mark.parametrize("use_torchaudio", [True, False], ids=["torchaudio", "restated"])
def test_getitem_restatement_at_other_sample_rates(golden_dir, use_torchaudio):
    """The unmodified reference with `torchaudio.load` reporting 8 ... 48 kHz: kaldi.fbank derives the frame geometry from the
    rate (kaldi.py:138-140); the oracle must follow (pins the generic-geometry kernel)."""
    g = _load(golden_dir, "getitem_rates.npz")
    conf = json.loads(str(g["conf"]))
    waves = tiny_dataset_waves()
    fill = ((torch.zeros(1) - conf["mean"]) / conf["std"]).numpy()[0]
    for sr in g["rates"]:
        sr = int(sr)
        shift, size, padded = O.frame_geometry(sr)
        assert (shift, size) == (int(sr * 0.010), int(sr * 0.025)) and padded >= size and padded & (padded - 1) == 0
        random.seed(sr)
        np.random.seed(sr)
        torch.manual_seed(sr)
        for i in range(len(waves)):
            out, _ = O.item_torch(waves, i, conf, random, np.random, use_torchaudio=use_torchaudio, sample_rate=sr)
            ref = g[f"feats_{sr}"][i]
            assert np.array_equal(ref == fill, out.numpy() == fill), (sr, i)
            assert np.abs(out.numpy() - ref).max() <= HOST_DRIFT, (sr, i)


def test_batch_oracle_equals_item_oracle():
    waves = tiny_dataset_waves()
    out = O.batch_torch(waves, [0, 5, 2], [-1, 1, 4], [0.0, 0.4, 0.6], [(0, 0, 0, 0), (3, 5, 10, 7), (100, 20, 90, 6)],
                        96, -4.6476, 4.5699)
    assert out.shape == (3, 96, 128)
    one = O.apply_masks_and_norm_torch(O.wav2fbank_torch(waves[5], waves[1], 0.4, 96), (3, 5, 10, 7), -4.6476, 4.5699)
    assert torch.equal(out[1], one)


def test_short_clip_raises():
    with pytest.raises(AssertionError):
        O.kaldi_fbank_np(np.zeros(399, np.float32))
    assert O.kaldi_fbank_np(np.zeros(400, np.float32), dtype=np.float32).shape == (1, 128)
    z = O.kaldi_fbank_np(np.zeros(1000, np.float32), dtype=np.float32)
    assert np.allclose(z, np.log(np.float32(O.EPS32)))


@pytest.mark.parametrize("use_torchaudio", [True, False])
def test_getitem_restatement_with_noise_and_roll(golden_dir, use_torchaudio):
    """audioset_dataset.py:197-199 (noise == True): every cell carries noise, so there is no fill pattern to compare;
    the values must agree with the unmodified reference's output under the same seeds (same three extra draws)."""
    g = _load(golden_dir, "getitem_noise.npz")
    conf = json.loads(str(g["conf"]))
    assert conf["noise"] is True
    waves = tiny_dataset_waves()
    k = 0
    for seed in g["seeds"]:
        random.seed(int(seed))
        np.random.seed(int(seed))
        torch.manual_seed(int(seed))
        for i in range(len(waves)):
            out, _ = O.item_torch(waves, i, conf, random, np.random, use_torchaudio=use_torchaudio)
            assert np.abs(out.numpy() - g["feats"][k]).max() <= HOST_DRIFT, k
            k += 1
    assert k == len(g["feats"])
    # the explicit form: roll by s moves row t to (t + s) mod T, noise is added before the roll
    fb, field = torch.arange(12.0).reshape(6, 2), torch.ones(6, 2)
    y = O.noise_roll_torch(fb, field, 0.5, -2)
    assert torch.allclose(y[4], fb[0] + 0.05) and torch.allclose(y[0], fb[2] + 0.05)
