"""CPU lane-stepping of the sm_100a streaming kernel's per-lane code (tests/emu/emu_tile.cc) against the oracle.

The harness compiles psla_b200/csrc/fbank_stream.cuh + fft_core.cuh with g++ and runs the kernel's phases one
simulated lane at a time.  It pins the index maps, twiddles, two-pass FFT, spectrum separation and the packed
mel layout where no GPU is available.  TEST ONLY (not loadable from psla_b200)."""
import ctypes
import os
import subprocess
import tempfile

import numpy as np
import pytest
import torch

from oracle import psla_oracle as O
from psla_b200.synth import synth_wave

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
F32P = ctypes.POINTER(ctypes.c_float)


def _p(a):
    return a.ctypes.data_as(F32P)


@pytest.fixture(scope="module")
def emu():
    out = os.path.join(tempfile.gettempdir(), f"psla_emu_{os.getuid()}.so")
    srcs = [os.path.join(ROOT, "tests", "emu", "emu_tile.cc"), os.path.join(ROOT, "psla_b200", "csrc", "tables.cc")]
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-ffp-contract=off", "-o", out, *srcs], check=True)
    return ctypes.CDLL(out)


@pytest.mark.parametrize("n", [16, 32, -32])
def test_register_fft_vs_direct_dft(emu, n):
    """n = -32: the pruned 32-point FFT whose inputs 25..31 are zero (400 = 25 * 16 samples per frame)."""
    N = abs(n)
    rng = np.random.default_rng(N)
    a, b = rng.standard_normal(N).astype(np.float32), rng.standard_normal(N).astype(np.float32)
    if n < 0:
        a[25:] = 0
        b[25:] = 0
    orr, oi = np.zeros(N, np.float32), np.zeros(N, np.float32)
    emu.emu_fft(n, _p(a), _p(b), _p(orr), _p(oi))
    ref = np.fft.fft(a.astype(np.float64) + 1j * b)
    assert np.abs(orr + 1j * oi - ref).max() < 3e-6


def _native_tables(emu, num_bins=128):
    win, mel, tab = np.zeros(400, np.float32), np.zeros((num_bins, 256), np.float32), np.zeros(8192, np.float32)
    spec = ctypes.c_int(0)
    n = emu.emu_tables(num_bins, _p(win), _p(mel), _p(tab), tab.size, ctypes.byref(spec))
    assert n > 0
    lay = (ctypes.c_int * 9)()
    emu.emu_layout(lay)
    return win, mel, tab[:n], bool(spec.value), list(lay)


def test_native_tables_vs_torch_order_tables(emu, golden_dir):
    win, mel, tab, spec, (o_win, o_tw, o_twz, o_ms, o_mb, o_mi, o_mw, win_row, ex_row) = _native_tables(emu)
    g = np.load(os.path.join(golden_dir, "tables.npz"))
    assert np.abs(win - g["hann400"]).max() <= 1.2e-7
    ref = O.mel_banks_torch().numpy()
    assert np.array_equal(mel != 0, ref != 0)
    assert np.abs(mel - ref).max() <= 2e-5  # libm logf vs torch's vectorised log: an ulp of mel(f) ~ 2840
    assert spec  # 16 kHz / 128 bins fits the unrolled (2, 3, 6, 10) mel path
    # transposed half window
    wt = tab[o_win:o_win + 16 * win_row].reshape(16, win_row)
    n2, n1 = np.meshgrid(np.arange(16), np.arange(25), indexing="ij")
    assert np.array_equal(wt[:, :25], 0.5 * win[16 * n1 + n2]) and not wt[:, 25:].any()
    # pass-2 twiddles: row t >= 1 is W_512^(t n2), row 0 is W_32^n2; TWZ is W_32^-n2
    tw = tab[o_tw:o_tw + 16 * ex_row].reshape(16, ex_row)[:, :32]
    t, n2 = np.meshgrid(np.arange(16), np.arange(16), indexing="ij")
    z = np.where(t == 0, np.exp(-2j * np.pi * n2 / 32), np.exp(-2j * np.pi * t * n2 / 512))
    assert np.abs(tw[:, 0::2] - z.real).max() < 6e-8 and np.abs(tw[:, 1::2] - z.imag).max() < 6e-8
    twz = tab[o_twz:o_twz + 32]
    zz = np.exp(2j * np.pi * np.arange(16) / 32)
    assert np.abs(twz[0::2] - zz.real).max() < 6e-8 and np.abs(twz[1::2] - zz.imag).max() < 6e-8
    # packed mel: rebuilding the dense matrix from (start, bin, weights) gives the dense table back
    start = tab[o_ms:o_ms + 128].view(np.int32).reshape(4, 32)
    mbin = tab[o_mb:o_mb + 128].view(np.int32).reshape(4, 32)
    info = tab[o_mi:o_mi + 8].view(np.int32)
    w = tab[o_mw:].reshape(-1, 32)[:info[:4].sum()]
    dense = np.zeros_like(mel)
    for i in range(4):
        assert sorted(mbin[i]) == list(range(32 * i, 32 * i + 32))  # a permutation inside the slot: one 128 B output line
        for q in range(4):  # the 8 lanes of a quarter-warp read 8 different bank groups (or the same bin)
            st = start[i, 8 * q:8 * q + 8]
            assert all(st[a] == st[c] or st[a] % 8 != st[c] % 8 for a in range(8) for c in range(a))
        for l in range(32):
            for r in range(info[i]):
                k = start[i, l] + r
                assert 0 <= k < 256
                dense[mbin[i, l], k] += w[info[4 + i] + r, l]
    assert np.array_equal(dense, mel)


@pytest.mark.parametrize("generic_mel", [0, 1])
@pytest.mark.parametrize("seed,n,kind", [(0, 160000, "gauss"), (7, 40000, "am"), (9, 12345, "int16"), (3, 400, "gauss"), (4, 719, "gauss")])
def test_emulated_kernel_vs_oracle(emu, seed, n, kind, generic_mel):
    if generic_mel and n > 40000:
        pytest.skip("one long clip is enough for the generic mel path")
    x = synth_wave(seed, n, kind)
    xm = (x - x.mean())
    m = O.num_frames(n)
    out = np.zeros((m, 128), np.float32)
    xin = np.ascontiguousarray(xm[0].numpy())
    mel = np.ascontiguousarray(O.mel_banks_torch().numpy())
    win = np.ascontiguousarray(torch.hann_window(400, periodic=False).numpy())
    assert emu.emu_fbank(_p(xin), n, 128, _p(mel), _p(win), generic_mel, _p(out)) == m
    ref32 = O.ref_fbank(xm).numpy()
    ref64 = O.ref_fbank(xm.double()).numpy()
    e_ref = np.abs(ref32 - ref64).max()
    assert np.abs(out - ref32).max() <= 1e-3
    assert np.abs(out - ref64).max() <= max(1.5 * e_ref, 2e-4)  # no worse than the fp32 reference itself
    assert out.shape == ref32.shape


def test_emulated_kernel_other_mel_sizes(emu):
    """num_mel_bins != 128 takes the run-time-loop mel path (tap counts differ from the unrolled (2, 3, 6, 10))."""
    x = synth_wave(11, 20000)
    xm = x - x.mean()
    for nb in (64, 40, 23):
        out = np.zeros((O.num_frames(20000), nb), np.float32)
        assert emu.emu_fbank(_p(np.ascontiguousarray(xm[0].numpy())), 20000, nb, None, None, 0, _p(out)) == out.shape[0]
        ref = O.ref_fbank(xm, num_mel_bins=nb).numpy()
        assert np.abs(out - ref).max() <= 1e-3


def test_emulated_kernel_zero_input_hits_the_floor(emu):
    out = np.ones((5, 128), np.float32)
    x = np.zeros(1040, np.float32)
    assert emu.emu_fbank(_p(x), 1040, 128, None, None, 0, _p(out)) == 5
    assert np.array_equal(out, np.full((5, 128), np.log(np.float32(O.EPS32)), np.float32))
    assert emu.emu_fbank(_p(x), 399, 128, None, None, 0, _p(out)) == -1


@pytest.mark.parametrize("generic_mel", [0, 1])
def test_phases_do_not_depend_on_the_lane_order(emu, generic_mel):
    """Race check without a GPU tool (compute-sanitizer is closed on the B200 pool): between two __syncwarp points the
    lanes of a warp run in no defined order, and three shared-memory regions alias each other (staged samples -> exchange
    buffer -> power spectra).  Stepping every phase in a seeded random lane order must give bit-identical features -- a
    lane reading inside a phase what another lane writes in the same phase would not."""
    x = synth_wave(11, 12000, "am")[0].numpy()
    x = np.ascontiguousarray(x - x.mean(dtype=np.float32))
    m = 1 + (len(x) - 400) // 160
    emu.emu_fbank_ordered.argtypes = [F32P, ctypes.c_int, ctypes.c_int, F32P, F32P, ctypes.c_int, ctypes.c_uint, F32P]
    outs = []
    for seed in (0, 1, 2, 3, 4):
        out = np.zeros((m, 128), np.float32)
        assert emu.emu_fbank_ordered(_p(x), len(x), 128, None, None, generic_mel, seed, _p(out)) == m
        outs.append(out)
    for o in outs[1:]:
        assert np.array_equal(o, outs[0])
