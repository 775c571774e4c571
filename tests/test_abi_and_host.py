"""CPU-only checks: the C-ABI library builds/loads and exports every symbol include/psla_frontend.h
declares; host-side logic (parameter draws in the reference's RNG order, packing, labels)."""
import json
import os
import random
import re
import tempfile

import numpy as np
import pytest
import torch

from psla_b200 import _lib, build, params as P, tables
from psla_b200.dataloaders import AudiosetDataset
from psla_b200.frontend import pack_waveforms
from psla_b200.synth import TINY_CLASSES, TINY_LABELS, tiny_dataset_waves

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "psla_frontend.h")).read()
    declared = set(re.findall(r"\b(psla_[a-z0-9_]+)\s*\(", header))
    declared -= {"psla_frontend_t"}
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.psla_version()
    assert lib.psla_strerror(-3).startswith(b"waveform shorter")


def test_no_cpu_fallback_in_product_package():
    """psla_b200 must not call the oracle, torchaudio's fbank, torch.fft or cuFFT: the CUDA library is the only path."""
    pkg = os.path.join(ROOT, "psla_b200")
    banned_py = ("import oracle", "from oracle", "kaldi.fbank(", "torch.fft", "compliance import", "FrequencyMasking(",
                 "TimeMasking(")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            path = os.path.join(dirpath, f)
            if f.endswith(".py"):
                code = [ln for ln in open(path).read().splitlines() if not ln.strip().startswith("#")]
                for ln in code:
                    assert not any(b in ln for b in banned_py), (f, ln)
            if f.endswith((".cu", ".cuh", ".cc", ".h")):
                src = open(path).read()
                assert "#include <cufft" not in src and "cufftExec" not in src, f


def test_tables_torch_order_match_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "tables.npz"))
    mel = tables.mel_banks().numpy()
    r, c = np.nonzero(mel)
    assert np.array_equal(r, g["mel_rows"]) and np.array_equal(c, g["mel_cols"])
    assert np.abs(mel[r, c] - g["mel_vals"]).max() <= 2e-5
    assert np.abs(tables.hann_window().numpy() - g["hann400"]).max() <= 1.2e-7


def test_mask_draw_matches_reference_modules(golden_dir):
    g = np.load(os.path.join(golden_dir, "mask_offsets.npz"))
    for s in range(200):
        torch.manual_seed(s)
        f0, fw, t0, tw = P.draw_masks(int(g["freqm"]), int(g["timem"]), int(g["F"]), int(g["T"]))
        assert (f0 if fw else 0, fw, t0 if tw else 0, tw) == tuple(int(v) for v in g["offsets"][s]), s
        assert 0 <= fw < 48 and 0 <= tw < 192 and f0 + fw <= 128 and t0 + tw <= 1056
    torch.manual_seed(0)
    before = torch.rand(1)
    torch.manual_seed(0)
    assert P.draw_masks(0, 0, 128, 1056) == (0, 0, 0, 0)  # nothing drawn for zero parameters (:182-185)
    assert torch.equal(torch.rand(1), before)


def _tiny_dataset(tmp, conf):
    waves = tiny_dataset_waves()
    store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
    jpath, cpath = os.path.join(tmp, "tiny.json"), os.path.join(tmp, "labels.csv")
    json.dump({"data": [{"wav": f"mem://{i}.wav", "labels": TINY_LABELS[i]} for i in range(len(waves))]}, open(jpath, "w"))
    with open(cpath, "w") as f:
        f.write("index,mid,display_name\n")
        for i, mid in enumerate(TINY_CLASSES):
            f.write(f"{i},{mid},class{i}\n")
    return AudiosetDataset(jpath, conf, cpath, loader=lambda p: (store[p].clone(), 16000), verbose=False), waves


@pytest.mark.parametrize("name", ["train", "eval", "skipnorm"])
def test_dataset_labels_and_rng_order_match_reference(golden_dir, name):
    """Same seeds -> same mixup decisions / lambdas (hence bit-identical soft labels) as the unmodified
    reference class, whose labels are stored in the golden file."""
    g = np.load(os.path.join(golden_dir, f"getitem_{name}.npz"))
    conf = json.loads(str(g["conf"]))
    with tempfile.TemporaryDirectory() as tmp:
        ds, waves = _tiny_dataset(tmp, conf)
        assert len(ds) == len(waves) and ds.label_num == 10
        k = 0
        for seed in g["seeds"]:
            random.seed(int(seed)); np.random.seed(int(seed)); torch.manual_seed(int(seed))
            for i in range(len(ds)):
                it = ds[i]
                assert np.array_equal(it.labels.numpy(), g["labels"][k]), (name, k)
                assert (it.wave2 is not None) == it.params.mixed
                # the fill pattern of the reference output reveals its mask offsets
                ref = g["feats"][k]
                fill = np.float32(0.0) if conf.get("skip_norm") else ((torch.zeros(1) - conf["mean"]) / conf["std"]).numpy()[0]
                f0, fw, t0, tw = it.params.mask
                cols = np.where((ref == fill).all(axis=0))[0]
                assert len(cols) == fw and (fw == 0 or cols[0] == f0), (name, k)
                if tw:
                    assert (ref[t0:t0 + tw] == fill).all()
                k += 1
        batch = ds.collate([ds[i] for i in range(4)], pinned=False)
        assert len(batch) == 4 and batch.labels.shape == (4, 10) and batch.masks.shape == (4, 4)
        assert (batch.waves.offsets % 4 == 0).all()


def test_pack_waveforms_alignment_and_content():
    ws = [torch.arange(5, dtype=torch.float32).unsqueeze(0), torch.ones(1, 8), torch.full((3,), 2.0)]
    p = pack_waveforms(ws)
    assert p.offsets.tolist() == [0, 8, 16] and p.lengths.tolist() == [5, 8, 3]
    assert torch.equal(p.wav[:5], ws[0][0]) and p.wav[5:8].abs().sum() == 0
    assert torch.equal(p.wav[16:19], ws[2])


def test_noise_draws_keep_the_reference_rng_order(golden_dir):
    """noise == True (audioset_dataset.py:197-199): three more draws per item (torch.rand(T, F), np.random.rand(),
    np.random.randint(-10, 10)).  numpy's generator also feeds the mixup lambdas, so bit-identical soft labels over two
    passes prove the draws are consumed in the reference's order; the fields are kept ("host") or dropped ("device")."""
    g = np.load(os.path.join(golden_dir, "getitem_noise.npz"))
    conf = json.loads(str(g["conf"]))
    with tempfile.TemporaryDirectory() as tmp:
        ds, waves = _tiny_dataset(tmp, conf)
        k = 0
        items = []
        for seed in g["seeds"]:
            random.seed(int(seed)); np.random.seed(int(seed)); torch.manual_seed(int(seed))
            for i in range(len(ds)):
                it = ds[i]
                assert np.array_equal(it.labels.numpy(), g["labels"][k]), k
                assert it.params.noise and 0.0 <= it.params.noise_scale < 1.0 and -10 <= it.params.roll < 10
                assert it.noise.shape == (96, 128) and 0.0 <= float(it.noise.min()) and float(it.noise.max()) < 1.0
                items.append(it)
                k += 1
        batch = ds.collate(items[:4], pinned=False)
        assert batch.noise.shape == (4, 96, 128) and batch.noise_scale.shape == (4,) and batch.roll.dtype == np.int32
        ds2, _ = _tiny_dataset(tmp, {**conf, "noise_field": "device"})
        random.seed(0); np.random.seed(0); torch.manual_seed(0)
        it2 = [ds2[i] for i in range(len(ds2))]
        assert all(a.noise is None for a in it2)
        assert [a.params.roll for a in it2] == [a.params.roll for a in items[:len(it2)]]  # same streams either way
        assert ds2.collate(it2, pinned=False).noise is None
        with pytest.raises(ValueError):
            _tiny_dataset(tmp, {**conf, "noise_field": "gpu"})


def test_frontend_needs_cuda():
    if torch.cuda.is_available():
        pytest.skip("has a GPU")
    from psla_b200.frontend import GpuFbankFrontend
    with pytest.raises(RuntimeError):
        GpuFbankFrontend({"num_mel_bins": 128, "target_length": 1056, "mean": 0.0, "std": 1.0})


def test_long_recording_chunk_offsets():
    from psla_b200.inference import chunk_offsets
    s, n = chunk_offsets(57_600_000, 160000)  # one hour at 16 kHz -> 360 windows of 10 s (BASELINE config 4)
    assert len(s) == 360 and (n == 160000).all() and s[-1] == 359 * 160000
    s, n = chunk_offsets(3 * 160000 + 7, 160000)  # a 7-sample remainder holds no frame: dropped (kaldi.py:142 would assert)
    assert s.tolist() == [0, 160000, 320000] and n.tolist() == [160000] * 3
    s, n = chunk_offsets(3 * 160000 + 500, 160000)  # a remainder with >= 1 frame is a short clip of its own
    assert s.tolist() == [0, 160000, 320000, 480000] and n.tolist() == [160000] * 3 + [500]
    s, n = chunk_offsets(500000, 160000, hop=80000)  # overlapping windows: only the first short one is kept
    assert s.tolist() == [0, 80000, 160000, 240000, 320000, 400000] and n[-1] == 100000
    assert len(chunk_offsets(399, 160000)[0]) == 0
    assert len(chunk_offsets(3 * 160000 + 500, 160000, keep_tail=False)[0]) == 3
    with pytest.raises(ValueError):
        chunk_offsets(1000, 100)


def test_bench_reference_arm_prints_one_json_line():
    """`bench.py --impl reference` (the CPU arm the driver launches next to ours): stdout is exactly one JSON line with
    the contract's keys; it runs the oracle port on the host cores, no GPU needed."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=root)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "clips/s" and d["higher_is_better"] is True and d["n_gpus"] == 1
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and "workload" in d["config"]


def test_norm_stats_accumulator():
    from psla_b200.dataloaders import NormStats, compute_norm_stats
    torch.manual_seed(0)
    xs = [torch.randn(3, 20, 8) * 4.5 - 4.6, torch.randn(2, 20, 8) * 4.5 - 4.6]
    mean, std = compute_norm_stats([(xs[0], None), xs[1]])
    allx = torch.cat([x.flatten() for x in xs]).double()
    assert abs(mean - allx.mean().item()) < 1e-9 and abs(std - allx.std(unbiased=False).item()) < 1e-9
    with pytest.raises(ValueError):
        NormStats().result()
