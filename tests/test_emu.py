"""CPU lane-stepping of the sm_100a tile kernel's per-lane code (tests/emu/emu_tile.cc) against the oracle.

The harness compiles psla_b200/csrc/fbank_tile.cuh + fft_core.cuh with g++ and runs the kernel's phases one
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


@pytest.mark.parametrize("n", [16, 32])
def test_register_fft_vs_direct_dft(emu, n):
    rng = np.random.default_rng(n)
    a, b = rng.standard_normal(n).astype(np.float32), rng.standard_normal(n).astype(np.float32)
    orr, oi = np.zeros(n, np.float32), np.zeros(n, np.float32)
    emu.emu_fft(n, _p(a), _p(b), _p(orr), _p(oi))
    ref = np.fft.fft(a.astype(np.float64) + 1j * b)
    assert np.abs(orr + 1j * oi - ref).max() < 3e-6


def test_native_tables_vs_torch_order_tables(emu, golden_dir):
    win, mel, tw = np.zeros(400, np.float32), np.zeros((128, 256), np.float32), np.zeros(1024, np.float32)
    emu.emu_tables(128, _p(win), _p(mel), _p(tw))
    g = np.load(os.path.join(golden_dir, "tables.npz"))
    assert np.abs(2 * win - g["hann400"]).max() <= 1.2e-7
    ref = O.mel_banks_torch().numpy()
    assert np.array_equal(mel != 0, ref != 0)
    assert np.abs(mel - ref).max() <= 2e-5  # libm logf vs torch's vectorised log: an ulp of mel(f) ~ 2840
    k1, n2 = np.meshgrid(np.arange(32), np.arange(16), indexing="ij")
    z = np.exp(-2j * np.pi * k1 * n2 / 512)
    assert np.abs(tw[0::2].reshape(32, 16) - z.real).max() < 6e-8 and np.abs(tw[1::2].reshape(32, 16) - z.imag).max() < 6e-8


@pytest.mark.parametrize("seed,n,kind", [(0, 160000, "gauss"), (7, 40000, "am"), (9, 12345, "int16"), (3, 400, "gauss"), (4, 719, "gauss")])
def test_emulated_kernel_vs_oracle(emu, seed, n, kind):
    x = synth_wave(seed, n, kind)
    xm = (x - x.mean())
    m = O.num_frames(n)
    out = np.zeros((m, 128), np.float32)
    xin = np.ascontiguousarray(xm[0].numpy())
    mel = np.ascontiguousarray(O.mel_banks_torch().numpy())
    win = np.ascontiguousarray(0.5 * torch.hann_window(400, periodic=False).numpy())
    assert emu.emu_fbank(_p(xin), n, 128, _p(mel), _p(win), _p(out)) == m
    ref32 = O.ref_fbank(xm).numpy()
    ref64 = O.ref_fbank(xm.double()).numpy()
    e_ref = np.abs(ref32 - ref64).max()
    assert np.abs(out - ref32).max() <= 1e-3
    assert np.abs(out - ref64).max() <= max(1.5 * e_ref, 2e-4)  # no worse than the fp32 reference itself
    assert out.shape == ref32.shape


def test_emulated_kernel_zero_input_hits_the_floor(emu):
    out = np.ones((5, 128), np.float32)
    x = np.zeros(1040, np.float32)
    assert emu.emu_fbank(_p(x), 1040, 128, None, None, _p(out)) == 5
    assert np.array_equal(out, np.full((5, 128), np.log(np.float32(O.EPS32)), np.float32))
    assert emu.emu_fbank(_p(x), 399, 128, None, None, _p(out)) == -1
