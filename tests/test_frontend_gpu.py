"""Parity of the CUDA frontend (through the C ABI) against the oracle on a B200.  `-m gpu`.

Tolerances (BASELINE.json north_star): <= 1e-3 absolute on log-mel values against the reference's fp32
torchaudio path on identical synthetic broadband waveforms; SpecAugment / padding cells bit-exact.
Features are compared in LOG-MEL units: normalised outputs are multiplied back by norm_div.

The fp32 reference is itself up to ~1.5e-3 away from an fp64 evaluation on a handful of near-floor cells per
10^7 (tests/test_oracle_golden.py, SURVEY.md 7 hard part 1), so on the large batch the 1e-3 bound is asserted for
every cell that is not within a factor ~20 of the log floor; the near-floor cells are bounded by 4e-3, their count
beyond 1e-3 is asserted to be a few per 10^7, and against fp64 truth we must be as accurate as the reference.
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
CONF = {"num_mel_bins": 128, "target_length": 1056, "mean": MEAN, "std": STD}
TOL = 1e-3


def _fill(mean=MEAN, div=STD):
    return ((torch.zeros(1) - mean) / div).item()


@pytest.fixture(scope="module")
def FE():
    from psla_b200 import build
    build.build()
    from psla_b200.frontend import GpuFbankFrontend
    return GpuFbankFrontend


@pytest.fixture(scope="module", params=[False, True], ids=["fused-mean", "exact-mean"])
def fe(FE, request):
    """Both treatments of the whole-clip DC removal (include/psla_frontend.h: psla_frontend_set_exact_clip_mean)."""
    return FE(CONF, exact_clip_mean=request.param)


def _same_fill_cells(out, ref, fill, div=STD, tol=TOL):
    """Masked / padded cells hold the same fp32 constant in both.  A LIVE cell may round to that constant by chance in
    one of the two (seen: 1 cell in 13.5 M); such a cell must still agree within the tolerance."""
    a, b = out == fill, ref == fill
    mism = a ^ b
    assert mism.sum().item() <= max(1, int(2e-7 * out.numel()))
    if mism.any():
        assert ((out - ref).abs()[mism] * div).max().item() <= tol


def _dead_cells(lengths, masks, T, F=128):
    """Cells the reference fills with 0.0 before normalisation: padding rows, time-mask rows, frequency-mask columns."""
    dead = torch.zeros(len(lengths), T, F, dtype=torch.bool)
    for b, (n, (f0, fw, t0, tw)) in enumerate(zip(lengths, masks)):
        dead[b, min(O.num_frames(n), T):] = True
        dead[b, t0:t0 + tw] = True
        dead[b, :, f0:f0 + fw] = True
    return dead


def _check(out, ref, div=STD, tol=TOL, fill=None, near_floor_below=None):
    """near_floor_below: log-mel value (for MEAN / div normalised outputs) at or below which a live cell counts as
    "near the float-epsilon floor" -- there the fp32 reference itself is > 1e-3 from fp64 truth (see the C2 test), so
    such cells (at most 0.2 % of the batch) are bounded by 4e-3 instead of tol."""
    out, ref = out.cpu(), ref.cpu()
    assert out.shape == ref.shape and out.dtype == torch.float32
    assert not torch.isnan(out).any()
    if fill is not None:  # mask placement and padding: same cells, same fp32 constant
        _same_fill_cells(out, ref, fill, div, tol)
    cell_err = (out - ref).abs() * div
    if near_floor_below is not None:
        at_floor = ref == ((torch.log(torch.tensor(O.EPS32, dtype=torch.float32)) - MEAN) / div)  # exactly log(eps): bit-identical
        # (a cell whose energy equals eps to within rounding may be clamped in one implementation and not in the other:
        # a handful per 10^7 cells, and those agree to within the tolerance)
        off = out[at_floor] != ref[at_floor]
        assert off.sum().item() <= max(0, int(1e-5 * at_floor.sum().item()))
        if off.any():
            assert ((out[at_floor] - ref[at_floor]).abs()[off] * div).max().item() <= tol
        near = (ref * div + MEAN <= near_floor_below) & (ref != (fill if fill is not None else float("nan"))) & ~at_floor
        assert near.sum().item() <= 2e-3 * ref.numel()
        assert cell_err[near].max().item() <= 4e-3 if near.any() else True
        cell_err = cell_err[~near]
    err = cell_err.max().item()
    assert err <= tol, err
    return err


@pytest.mark.parametrize("fast_log", [False, True])
@pytest.mark.parametrize("tables_from", ["torch", "native"])
def test_c1_single_clip(FE, golden_dir, fast_log, tables_from):
    fe = FE(CONF, fast_log=fast_log, tables_from=tables_from)
    x = synth_wave(0, 160000)
    out = fe([x])
    assert out.is_cuda and out.shape == (1, 1056, 128)
    ref = O.batch_torch([x], [0], [-1], [0.0], [(0, 0, 0, 0)], 1056, MEAN, STD)
    _check(out, ref, fill=_fill())
    assert (out[0, 998:].cpu() == _fill()).all()
    # committed golden rows of the unmodified reference (un-normalised log-mel)
    g = np.load(os.path.join(golden_dir, "fbank_c1.npz"))
    raw = out[0].cpu().double() * STD + MEAN
    assert np.abs(raw[g["rows"]].numpy() - g["fb32_rows"]).max() <= TOL
    # vs fp64 truth we are no worse than the fp32 reference is
    ref64 = O.batch_torch([x], [0], [-1], [0.0], [(0, 0, 0, 0)], 1056, MEAN, STD, dtype=torch.float64)
    d_ours, d_ref = out.cpu().double() - ref64, ref.double() - ref64
    # rms: the stable statistic; the max over 135k cells sits on a near-floor cell and moves with any change of rounding order
    assert d_ours.pow(2).mean().sqrt().item() <= 1.5 * d_ref.pow(2).mean().sqrt().item()
    assert d_ours.abs().max().item() <= 2.0 * d_ref.abs().max().item()
    fe.close()


def test_mixup_masks_and_corner_cases(fe):
    waves = [synth_wave(1, 160000), synth_wave(2, 160000, "am"), synth_wave(3, 100000, "int16"),
             synth_wave(4, 200000), synth_wave(5, 5000), synth_wave(6, 400), synth_wave(7, 163)]
    #        plain  equal-len  partner longer(cut)  partner shorter(pad)  self-mix  short  1-frame  tiny partner
    src1 = [0, 1, 2, 0, 2, 4, 5, 0]
    src2 = [-1, 0, 3, 2, 2, -1, -1, 6]
    lam = [0.0, 0.3, 0.55, 0.7, 0.5, 0.0, 0.0, 0.41]
    masks = [(0, 0, 0, 0), (10, 20, 100, 150), (0, 47, 900, 156), (127, 1, 0, 1), (50, 0, 500, 0), (3, 5, 2, 3),
             (0, 128, 0, 0), (100, 28, 1055, 1)]
    out = fe(waves, src1, src2, lam, masks)
    ref = O.batch_torch(waves, src1, src2, lam, masks, 1056, MEAN, STD)
    _check(out, ref, fill=_fill(), near_floor_below=-13.0)  # the "am" clip has near-silent frames
    assert (out[6].cpu() == _fill()).all()  # fully frequency-masked clip


def test_ragged_fsd50k_shape(FE):
    """C3: variable lengths up to > 30 s, padded / truncated to target_length = 3000, mixup across lengths."""
    fe = FE({**CONF, "target_length": 3000})
    rng = np.random.default_rng(1)
    lens = [480000, 500000, 130000] + [int(v) for v in rng.integers(4800, 480000, 5)]
    waves = [synth_wave(20 + i, n, ("gauss", "am", "int16")[i % 3]) for i, n in enumerate(lens)]
    B = len(waves)
    src1 = list(range(B))
    src2 = [1, -1, 0, 2, -1, 6, 5, -1]
    lam = [0.5, 0, 0.35, 0.62, 0, 0.48, 0.51, 0]
    torch.manual_seed(3)
    masks = [O.draw_mask_offsets(48, 192, 128, 3000) for _ in range(B)]
    out = fe(waves, src1, src2, lam, masks)
    ref = O.batch_torch(waves, src1, src2, lam, masks, 3000, MEAN, STD)
    _check(out, ref, fill=_fill(), near_floor_below=-13.0)  # "am" clips have near-silent frames
    fe.close()


@pytest.mark.parametrize("B", [24, 96])
def test_c3_fsd50k_batches_at_recipe_size(FE, B):
    """C3 at the reference's own batch size (egs/fsd50k/run.sh:21 batch_size=24; 96 = the sweep's next size) and recipe
    (:24-29 freqm 48, timem 192, mixup 0.5; :60 target_length 3000): clip lengths 0.3 .. 30 s, every other clip mixed with a
    partner of any length (shorter partners need the statistics pass), int16 PCM input (what the files hold)."""
    from psla_b200.frontend import to_pcm16
    fe = FE({**CONF, "target_length": 3000})
    rng = np.random.default_rng(B)
    lens = [int(v) for v in rng.integers(4800, 480001, B)]
    lens[0], lens[1] = 480000, 400
    waves = [to_pcm16(synth_wave(700 + i, n, ("gauss", "am", "int16")[i % 3])) for i, n in enumerate(lens)]
    src1 = list(range(B))
    src2 = [int(rng.integers(0, B)) if rng.random() < 0.5 else -1 for _ in range(B)]
    lam = [float(rng.beta(10, 10)) if s >= 0 else 0.0 for s in src2]
    torch.manual_seed(B)
    masks = [O.draw_mask_offsets(48, 192, 128, 3000) for _ in range(B)]
    out = fe(waves, src1, src2, lam, masks)
    assert sum(1 for a, c in zip(src1, src2) if c >= 0 and lens[c] < lens[a]) >= 2  # the statistics kernel ran
    torch.set_num_threads(max(1, min(16, os.cpu_count() or 1)))
    ref = O.batch_torch([w.float() / 32768.0 for w in waves], src1, src2, lam, masks, 3000, MEAN, STD)
    _check(out, ref, fill=_fill(), near_floor_below=-13.0)
    fe.close()


def test_truncation_and_odd_target_length(FE):
    fe = FE({**CONF, "target_length": 77})
    waves = [synth_wave(31, 160000), synth_wave(32, 9000)]
    out = fe(waves, [0, 1], [-1, 0], [0, 0.6], [(0, 0, 70, 7), (5, 5, 0, 0)])
    ref = O.batch_torch(waves, [0, 1], [-1, 0], [0, 0.6], [(0, 0, 70, 7), (5, 5, 0, 0)], 77, MEAN, STD)
    _check(out, ref, fill=_fill())
    fe.close()


def test_skip_norm_and_std_scale(FE):
    waves = [synth_wave(41, 50000), synth_wave(42, 30000)]
    args = ([0, 1], [1, -1], [0.45, 0], [(3, 10, 100, 50), (0, 0, 0, 0)])
    fe = FE({**CONF, "target_length": 320, "skip_norm": True})
    out = fe(waves, *args)
    ref = O.batch_torch(waves, *args, 320, MEAN, STD, skip_norm=True)
    _check(out, ref, div=1.0, fill=0.0)
    fe.close()
    fe2 = FE({**CONF, "target_length": 320}, norm_std_scale=2.0)  # the `/ (norm_std * 2)` form of the north_star text
    out2 = fe2(waves, *args)
    ref2 = O.batch_torch(waves, *args, 320, MEAN, STD * 2)
    _check(out2, ref2, div=2 * STD, fill=_fill(MEAN, STD * 2))
    fe2.close()


def test_zero_clip_and_floor(fe):
    z = torch.zeros(1, 16000)
    out = fe([z], target_length=128)
    ref = O.batch_torch([z], [0], [-1], [0.0], [(0, 0, 0, 0)], 128, MEAN, STD)
    assert torch.equal(out.cpu(), ref)  # every live cell is log(eps), normalised: bit-exact
    for fast in (False, True):
        fe.set_fast_log(fast)
        assert torch.equal(fe([z], target_length=128).cpu(), ref)
    fe.set_fast_log(True)


def test_errors_mirror_the_reference(fe):
    from psla_b200._lib import PslaError, ShortClipError
    with pytest.raises(AssertionError):  # kaldi.py:142 raises AssertionError for < 400 samples
        fe([synth_wave(1, 399)])
    with pytest.raises(ShortClipError):
        fe([synth_wave(1, 16000), synth_wave(2, 10)], [0, 1])
    ok = [synth_wave(1, 16000)]
    with pytest.raises(PslaError):
        fe(ok, [0], masks=[(120, 20, 0, 0)])
    with pytest.raises(PslaError):
        fe(ok, [0], masks=[(0, 0, 1000, 100)])
    with pytest.raises(PslaError):
        fe(ok, [1])
    with pytest.raises(PslaError):
        fe(ok, [0], [3], [0.5])
    assert fe(ok, []).shape == (0, 1056, 128)
    # a descriptor that reaches past the waveform buffer is an error of the call, not an out-of-bounds device read
    from psla_b200.frontend import PackedWaves
    buf = synth_wave(1, 16000)[0].cuda()
    for offs, lens in (([0], [16001]), ([15000], [2000]), ([16000], [400]), ([-4], [4000])):
        with pytest.raises(PslaError):
            fe.forward_packed(PackedWaves(buf, np.array(offs, np.int64), np.array(lens, np.int64)), [0])
    with pytest.raises(PslaError):
        fe.forward_host(PackedWaves(buf.cpu(), np.array([8000], np.int64), np.array([9000], np.int64)), [0])
    assert fe.forward_packed(PackedWaves(buf, np.array([15600], np.int64), np.array([400], np.int64)), [0]).shape == (1, 1056, 128)


def test_guard_bands_around_every_buffer(fe):
    """Memory-safety evidence without a tool (compute-sanitizer is closed on this pool): waveforms of awkward lengths at
    aligned and unaligned offsets are surrounded by NaN (fp32) / full-scale (int16) guard samples, the output is a window of
    a larger sentinel-filled buffer.  Features equal the oracle's, so no guard sample was ever used (a NaN would spread
    over its whole frame), and no byte outside the output window was written -- for every descriptor path and layout."""
    from psla_b200.frontend import PackedWaves, to_pcm16
    T = 300
    lens = [400, 401, 559, 560, 561, 1039, 12345, 48000, 47999, 16003]
    gaps = [0, 1, 2, 3, 4, 5, 7, 9, 13, 64]  # guard samples in front of each waveform: starts at every alignment
    waves = [synth_wave(500 + i, n, ("gauss", "int16")[i % 2]) for i, n in enumerate(lens)]
    src1 = list(range(10)) + [7, 8]
    src2 = [-1, 0, -1, 2, -1, 6, 7, 8, 7, -1, 9, 6]  # longer, shorter and equal partners
    lam = [0.0, 0.4, 0.0, 0.6, 0.0, 0.5, 0.3, 0.7, 0.45, 0.0, 0.55, 0.35]
    torch.manual_seed(3)
    masks = [O.draw_mask_offsets(48, 100, 128, T) for _ in src1]
    ref = O.batch_torch(waves, src1, src2, lam, masks, T, MEAN, STD)
    ref16 = O.batch_torch([to_pcm16(w).float() / 32768.0 for w in waves], src1, src2, lam, masks, T, MEAN, STD)
    for pcm in (False, True):
        offs, pos = [], 32
        for g_, n in zip(gaps, lens):
            pos += g_
            offs.append(pos)
            pos += n
        total = pos + 32
        if pcm:
            flat = torch.full((total,), 32767, dtype=torch.int16)
            for o, w in zip(offs, waves):
                flat[o:o + w.shape[-1]] = to_pcm16(w)[0]
        else:
            flat = torch.full((total,), float("nan"))
            for o, w in zip(offs, waves):
                flat[o:o + w.shape[-1]] = w[0]
        pk = PackedWaves(flat.cuda(), np.array(offs, np.int64), np.array(lens, np.int64))
        for layout in (("tf", "ft", "ft_bf16") if not fe.exact_clip_mean else ("tf",)):
            fe.set_output_layout(layout)
            try:
                for pmeta in (True, False):
                    fe.set_param_meta(pmeta)
                    shape, dtype = fe.out_shape_dtype(len(src1), T)
                    n_out = int(np.prod(shape))
                    sentinel = -12345.0 if dtype == torch.float32 else -12352.0
                    big = torch.full((n_out + 2 * 4096,), sentinel, dtype=dtype, device=fe.device)
                    out = big[4096:4096 + n_out].view(shape)
                    fe.forward_packed(pk, src1, src2, lam, masks, out=out, target_length=T)
                    torch.cuda.synchronize()
                    assert (big[:4096] == sentinel).all() and (big[4096 + n_out:] == sentinel).all()
                    got = out.float().cpu() if layout == "tf" else out.float().cpu().transpose(1, 2)
                    assert not torch.isnan(got).any()
                    want = ref16 if pcm else ref
                    tol = TOL if layout != "ft_bf16" else 0.03  # bf16 rounding of values up to ~4
                    live = want != _fill()
                    assert ((got - want).abs() * STD)[live].max().item() <= (4e-3 if layout != "ft_bf16" else 0.1) and \
                        ((got - want).abs() * STD)[live].median().item() <= tol
            finally:
                fe.set_output_layout("tf")
                fe.set_param_meta(True)


def test_outlier_samples_do_not_raise_the_noise_floor(fe):
    """the default clip-mean mode subtracts a per-clip constant instead of the mean.  With the clip's FIRST
    sample as that constant, a full-scale click at sample 0 in front of a quiet 16-bit body put every staged value at
    magnitude 0.9 and 1 % of the cells were off by up to 1.2e-2 (mixed clips: lambda * x rounds at ulp(0.9)).  The
    constant is now a five-sample median of the clip; click, clipped first samples and a loud onset must all stay within
    the tolerance, mixed and unmixed, fp32 and int16."""
    from psla_b200.frontend import to_pcm16

    def body(seed, n, amp):  # low-pass noise: nothing but quantisation noise above ~1 kHz
        g = torch.Generator().manual_seed(seed)
        x = torch.randn(n + 64, generator=g, dtype=torch.float64)
        k = torch.hann_window(65, periodic=False, dtype=torch.float64)
        y = torch.nn.functional.conv1d(x.view(1, 1, -1), (k / k.sum()).view(1, 1, -1)).view(-1)[:n]
        return (amp * y / y.std()).float().unsqueeze(0)

    n = 80000
    waves = []
    for i in range(6):
        x = body(40 + i, n, 4e-3)
        if i % 3 == 0:
            x[0, 0] = 0.9  # click at the very first sample
        elif i % 3 == 1:
            x[0, :3] = -1.0  # three clipped samples
        else:
            x[0, :1500] += 0.5 * synth_wave(60 + i, 1500, "gauss", amp=0.6)[0].clamp(-1, 1)  # loud onset
        waves.append(x)
    src1, src2 = list(range(6)), [1, -1, 3, 0, -1, 4]
    lam = [0.37, 0.0, 0.61, 0.43, 0.0, 0.29]
    masks = [(0, 0, 0, 0)] * 6
    for pcm in (False, True):
        ws = [to_pcm16(w) for w in waves] if pcm else waves
        ref_in = [w.float() / 32768.0 for w in ws] if pcm else waves
        out = fe(ws, src1, src2, lam, masks, target_length=500)
        ref = O.batch_torch(ref_in, src1, src2, lam, masks, 500, MEAN, STD)
        ref64 = O.batch_torch(ref_in, src1, src2, lam, masks, 500, MEAN, STD, dtype=torch.float64)
        live = ref != _fill()
        err = ((out.cpu() - ref).abs() * STD)[live]
        ref_err = ((ref.double() - ref64).abs() * STD)[live]
        assert (err > TOL).sum().item() <= max(3, 2 * int((ref_err > TOL).sum())), ((err > TOL).sum().item(), err.max().item())
        assert err.max().item() <= max(2e-3, 1.5 * ref_err.max().item())


def test_unaligned_offsets_and_shared_buffer(fe):
    """C4-style chunking: clips are windows of one long recording at arbitrary (also non multiple-of-4) offsets."""
    from psla_b200.frontend import PackedWaves
    long = synth_wave(50, 160000 * 3 + 7)
    offs = np.array([0, 160000, 320003, 1], dtype=np.int64)
    lens = np.array([160000, 160000, 160000, 160001], dtype=np.int64)
    packed = PackedWaves(long[0].cuda(), offs, lens)
    out = fe.forward_packed(packed, [0, 1, 2, 3], [-1, 2, -1, 0], [0, 0.5, 0, 0.3])
    waves = [long[:, o:o + n] for o, n in zip(offs, lens)]
    ref = O.batch_torch(waves, [0, 1, 2, 3], [-1, 2, -1, 0], [0, 0.5, 0, 0.3], [(0, 0, 0, 0)] * 4, 1056, MEAN, STD)
    _check(out, ref, fill=_fill())


def test_host_entry_point_equals_device_entry_point(fe):
    from psla_b200.frontend import pack_waveforms
    waves = [synth_wave(60 + i, 160000) for i in range(4)]
    args = ([0, 1, 2], [3, -1, 0], [0.4, 0, 0.6], [(1, 2, 3, 4), (0, 0, 0, 0), (100, 20, 900, 100)])
    dev = fe(waves, *args).cpu()
    host_in = pack_waveforms(waves, pinned=True)
    out_host = torch.empty(3, 1056, 128, pin_memory=True)
    fe.forward_host(host_in, *args, out_host=out_host)
    torch.cuda.synchronize()
    assert torch.equal(out_host, dev)
    view = fe.forward_host(host_in, *args)
    torch.cuda.synchronize()
    assert view.is_cuda and torch.equal(view.cpu(), dev)


def test_deterministic_and_batch_independent(fe):
    waves = [synth_wave(70 + i, 160000) for i in range(6)]
    a = fe(waves, [0, 1, 2, 3], [4, -1, 5, -1], [0.3, 0, 0.8, 0]).cpu()
    b = fe(waves, [0, 1, 2, 3], [4, -1, 5, -1], [0.3, 0, 0.8, 0]).cpu()
    assert torch.equal(a, b)
    c = fe(waves, [2], [5], [0.8]).cpu()
    assert torch.equal(c[0], a[2])


def test_pcm16_input_equals_decoded_float_input(fe):
    """16-bit PCM as it sits in a wav file -> same bits as the fp32 entry point fed with the decoder's floats
    (torchaudio.load: value / 32768; audioset_dataset.py:91,95-96), on the device and the host entry points, and the
    reference's fp32 path on those floats within the tolerance."""
    from psla_b200.frontend import pack_waveforms, to_pcm16
    lens = [160000, 100000, 200000, 5003, 400, 48001]
    pcm = [to_pcm16(synth_wave(500 + i, n, ("gauss", "am", "int16")[i % 3])) for i, n in enumerate(lens)]
    flt = [w.to(torch.float32) / 32768.0 for w in pcm]
    src1 = [0, 1, 2, 0, 3, 4, 5, 1]
    src2 = [-1, 0, 1, 2, -1, -1, 0, 5]   # equal / longer / shorter partners (the latter need the statistics pass)
    lam = [0.0, 0.3, 0.55, 0.7, 0.0, 0.0, 0.45, 0.62]
    masks = [(0, 0, 0, 0), (10, 20, 100, 150), (0, 47, 900, 156), (127, 1, 0, 1), (50, 0, 500, 0), (3, 5, 2, 3), (0, 0, 0, 0),
             (100, 28, 1055, 1)]
    a = fe(pcm, src1, src2, lam, masks).cpu()
    b = fe(flt, src1, src2, lam, masks).cpu()
    assert torch.equal(a, b)
    ref = O.batch_torch(flt, src1, src2, lam, masks, 1056, MEAN, STD)
    _check(a, ref, fill=_fill(), near_floor_below=-13.0)  # "am" clips have near-silent frames
    host = pack_waveforms(pcm, pinned=True)
    assert host.wav.dtype == torch.int16
    out_host = torch.empty(len(src1), 1056, 128, pin_memory=True)
    fe.forward_host(host, np.array(src1, np.int32), np.array(src2, np.int32), np.array(lam), np.array(masks, np.int32), out_host=out_host)
    torch.cuda.synchronize()
    assert torch.equal(out_host, a)
    # unaligned starts inside one shared buffer (scalar load path)
    flat = torch.cat([torch.zeros(3, dtype=torch.int16), pcm[0][0], torch.zeros(2, dtype=torch.int16), pcm[1][0]]).to(fe.device)
    from psla_b200.frontend import PackedWaves
    pk = PackedWaves(flat, np.array([3, 3 + lens[0] + 2], np.int64), np.array(lens[:2], np.int64))
    c = fe.forward_packed(pk, np.array([0, 1], np.int32), np.array([-1, 0], np.int32), np.array([0.0, 0.3])).cpu()
    d = fe(flt[:2], [0, 1], [-1, 0], [0.0, 0.3]).cpu()
    assert torch.equal(c, d)


def test_launch_plumbing_does_not_change_results(fe):
    """Descriptors as kernel parameters vs uploaded, launch overlap, copy stream, cost model: identical bits.  A batch
    above the 128-clip parameter-space limit (upload path) equals the same clips computed in small batches."""
    from psla_b200 import params as P
    from psla_b200.frontend import pack_waveforms
    torch.manual_seed(5)
    waves = [synth_wave(300 + i, 16000 + 977 * i) for i in range(12)]
    packed = pack_waveforms(waves, device=fe.device)
    B = 150
    src1 = np.arange(B, dtype=np.int32) % 12
    src2 = np.where(np.arange(B) % 3 == 0, (np.arange(B) + 5) % 12, -1).astype(np.int32)
    lam = np.random.default_rng(3).beta(10, 10, B)
    masks = np.array([P.draw_masks(48, 60, 128, 200) for _ in range(B)], dtype=np.int32)
    big = fe.forward_packed(packed, src1, src2, lam, masks, target_length=200).cpu()
    parts = [fe.forward_packed(packed, src1[i:i + 50], src2[i:i + 50], lam[i:i + 50], masks[i:i + 50], target_length=200).cpu()
             for i in range(0, B, 50)]
    assert torch.equal(big, torch.cat(parts))
    # clips whose partner is longer need no statistics pass (fused-mean mode): these go through parameter space
    n = 50
    s1 = (np.arange(n) % 6).astype(np.int32)
    s2 = np.where(np.arange(n) % 3 == 0, 6 + np.arange(n) % 6, -1).astype(np.int32)
    args = (packed, s1, s2, lam[:n], masks[:n])
    ref = fe.forward_packed(*args, target_length=200).cpu()
    try:
        for setter, val, back in ((fe.set_param_meta, False, True), (fe.set_launch_overlap, "off", "stream"),
                                  (fe.set_launch_overlap, "call", "stream"), (fe.set_meta_stream, False, True)):
            setter(val)
            outs = [torch.empty(n, 200, 128, device=fe.device) for _ in range(3)]
            for o in outs:  # back-to-back launches into different buffers (what the overlap path keys on)
                fe.forward_packed(*args, out=o, target_length=200)
            torch.cuda.synchronize()
            assert all(torch.equal(o.cpu(), ref) for o in outs)
            setter(back)
        # from here on consecutive calls may overlap (PSLA_OVERLAP_STREAM: nothing but our own launches on the stream)
        outs = [torch.empty(n, 200, 128, device=fe.device) for _ in range(4)]
        for rep in range(3):
            for o in outs:
                fe.forward_packed(*args, out=o, target_length=200)
            fe.forward_packed(*args, out=outs[3], target_length=200)  # same buffer twice in a row: no overlap allowed
        torch.cuda.synchronize()
        assert all(torch.equal(o.cpu(), ref) for o in outs)
        # two buffers used alternately by tiny batches of DIFFERENT clips: the last writer of each buffer must win (a launch
        # may overlap its predecessor, never an earlier writer of the same range)
        singles = [fe.forward_packed(packed, s1[i:i + 1], None, None, None, target_length=200).cpu() for i in range(6)]
        bufs = [torch.empty(1, 200, 128, device=fe.device) for _ in range(2)]
        last = [None, None]
        for k in range(60):
            i = (k * 5 + k // 7) % 6
            fe.forward_packed(packed, s1[i:i + 1], None, None, None, out=bufs[k % 2], target_length=200)
            last[k % 2] = i
        torch.cuda.synchronize()
        assert torch.equal(bufs[0].cpu(), singles[last[0]]) and torch.equal(bufs[1].cpu(), singles[last[1]])
        fe.set_cost_model(32, 1, 0)
        assert torch.equal(fe.forward_packed(*args, target_length=200).cpu(), ref)
    finally:
        fe.set_param_meta(True)
        fe.set_launch_overlap("call")
        fe.set_meta_stream(True)
        fe.set_cost_model(64, 2, 48)


def test_overlapping_launches_keep_stream_order(FE):
    """PSLA_OVERLAP_STREAM (include/psla_frontend.h): a launch may START while its predecessor still runs, but it joins
    that predecessor before it exits, so (a) whatever is enqueued next sees every earlier launch finished -- checked with
    one heavy launch followed by tiny ones into rotating and aliasing buffers and an immediate consumer -- and (b) a
    plain kernel enqueued between two calls (here: torch kernels that REWRITE the waveform buffer) is waited for in full."""
    fe = FE({"num_mel_bins": 128, "target_length": 400, "mean": MEAN, "std": STD})
    fe.set_launch_overlap("stream")
    dev = fe.device
    from psla_b200.frontend import PackedWaves, pack_waveforms
    T = 400
    heavy_waves = [synth_wave(100 + s, 64000, ("gauss", "am")[s % 2]) for s in range(96)]
    tiny_waves = [synth_wave(300 + s, 8000 + 160 * s) for s in range(6)]
    heavy, tiny = pack_waveforms(heavy_waves, device=dev), pack_waveforms(tiny_waves, device=dev)
    s_heavy = np.arange(96, dtype=np.int32)
    fe.set_launch_overlap("off")
    ref_heavy = fe.forward_packed(heavy, s_heavy, target_length=T).clone()
    ref_tiny = [fe.forward_packed(tiny, np.array([i], np.int32), target_length=T).clone() for i in range(6)]
    torch.cuda.synchronize()
    fe.set_launch_overlap("stream")
    # (a) heavy, then 6 tiny launches: rotating over 3 buffers (every buffer written twice), consumer right behind
    out_heavy = torch.empty(96, T, 128, device=dev)
    bufs = [torch.empty(1, T, 128, device=dev) for _ in range(3)]
    for rep in range(40):
        out_heavy.zero_()
        for b_ in bufs:
            b_.zero_()
        fe.forward_packed(heavy, s_heavy, out=out_heavy, target_length=T)
        for k in range(6):
            fe.forward_packed(tiny, np.array([k], np.int32), out=bufs[k % 3], target_length=T)
        got_heavy = out_heavy.clone()  # a consumer kernel directly behind the tiny launches: the heavy one must be done
        got = [b_.clone() for b_ in bufs]
        torch.cuda.synchronize()
        assert torch.equal(got_heavy, ref_heavy), rep
        for j in range(3):
            assert torch.equal(got[j], ref_tiny[3 + j]), (rep, j)  # the LAST writer of each buffer
    # ... and aliasing: tiny launches writing INTO the heavy launch's output range go out as plain launches
    for rep in range(10):
        fe.forward_packed(heavy, s_heavy, out=out_heavy, target_length=T)
        for k in range(4):
            fe.forward_packed(tiny, np.array([k], np.int32), out=out_heavy[10 * k:10 * k + 1], target_length=T)
        torch.cuda.synchronize()
        expect = ref_heavy.clone()
        for k in range(4):
            expect[10 * k] = ref_tiny[k][0]
        assert torch.equal(out_heavy, expect), rep
    # (b) a producer kernel between two calls rewrites the waveform buffer the second call reads
    n = 64000
    a_src = torch.stack([w.reshape(-1) for w in heavy_waves[:48]]).to(dev)
    b_src = torch.stack([w.reshape(-1) for w in heavy_waves[48:]]).to(dev)
    offs, lens = np.arange(48, dtype=np.int64) * n, np.full(48, n, dtype=np.int64)
    s48 = np.arange(48, dtype=np.int32)
    buf = torch.empty(48 * n, device=dev)
    pk = PackedWaves(buf, offs, lens)
    o1, o2 = torch.empty(48, T, 128, device=dev), torch.empty(48, T, 128, device=dev)
    for rep in range(40):
        buf.view(48, n).copy_(b_src if rep % 2 else a_src)
        torch.cuda.synchronize()
        first, second = (b_src, a_src) if rep % 2 else (a_src, b_src)
        fe.forward_packed(pk, s48, out=o1, target_length=T)
        torch.add(second, 0.0, out=buf.view(48, n))  # an elementwise KERNEL (not a memcpy) overwriting the input
        fe.forward_packed(pk, s48, out=o2, target_length=T)
        torch.cuda.synchronize()
        r1, r2 = (ref_heavy[48:], ref_heavy[:48]) if rep % 2 else (ref_heavy[:48], ref_heavy[48:])
        assert torch.equal(o1, r1) and torch.equal(o2, r2), rep
    # (c) batches that need the statistics pass (shorter mixup partners): statistics kernel and streaming kernel are one
    # chain -- the statistics pass of call k + 1 may run behind the streaming kernel of call k (own partial-sum buffer),
    # the streaming kernel joins it only where a clip needs the sums.  Different batches back to back, rotating buffers.
    rag = [synth_wave(900 + i, n) for i, n in enumerate((64000, 20000, 48000, 9000, 33000, 64000, 5000, 40000))]
    pkr = pack_waveforms(rag, device=dev)
    batches = []
    for k in range(4):
        s1 = np.array([(k + i) % 8 for i in range(6)], np.int32)
        s2 = np.array([(k + 3 * i + 1) % 8 if i % 2 == 0 else -1 for i in range(6)], np.int32)
        lam = np.full(6, 0.3 + 0.1 * k)
        batches.append((s1, s2, lam))
    fe.set_launch_overlap("off")
    want = [fe.forward_packed(pkr, s1, s2, lam, None, target_length=T).clone() for s1, s2, lam in batches]
    torch.cuda.synchronize()
    for mode in ("call", "stream"):
        fe.set_launch_overlap(mode)
        outs = [torch.empty(6, T, 128, device=dev) for _ in range(3)]
        for rep in range(30):
            got = []
            for k, (s1, s2, lam) in enumerate(batches):
                o = outs[(rep + k) % 3]
                fe.forward_packed(pkr, s1, s2, lam, None, out=o, target_length=T)
                got.append(o.clone())
            torch.cuda.synchronize()
            assert all(torch.equal(g_, w_) for g_, w_ in zip(got, want)), (mode, rep)


def test_overlap_fuzz_random_batches(FE):
    """Random batches (1..128 clips, ragged lengths, mixup with shorter / longer partners, masks, three target lengths) back
    to back in PSLA_OVERLAP_STREAM mode into output ranges carved from ONE arena (so consecutive ranges sometimes overlap
    and the library has to fall back to a plain launch), each result cloned by a consumer right behind the call:
    bit-identical to the same calls with overlap off.  Part of the substitute for compute-sanitizer (profiles/r02a/sanitizer.md)."""
    from psla_b200.frontend import pack_waveforms
    fe = FE({"num_mel_bins": 128, "target_length": 256, "mean": MEAN, "std": STD})
    dev = fe.device
    rng = np.random.default_rng(2024)
    pool = [synth_wave(4000 + i, int(rng.integers(400, 48000)), ("gauss", "am", "int16")[i % 3]) for i in range(40)]
    pk = pack_waveforms(pool, device=dev)
    calls = []
    for _ in range(48):
        B = int(rng.choice([1, 2, 3, 7, 24, 60, 128]))
        T = int(rng.choice([64, 256, 301]))
        s1 = rng.integers(0, 40, B).astype(np.int32)
        s2 = np.where(rng.random(B) < 0.5, rng.integers(0, 40, B), -1).astype(np.int32)
        lam = rng.beta(10, 10, B)
        masks = np.zeros((B, 4), np.int32)
        for b in range(B):
            if rng.random() < 0.5:
                masks[b] = (int(rng.integers(0, 100)), int(rng.integers(0, 28)), int(rng.integers(0, T - 40)), int(rng.integers(0, 40)))
        calls.append((B, T, s1, s2, lam, masks))
    arena = torch.empty(3 * 128 * 301 * 128, device=dev)
    starts = [int(rng.integers(0, 2 * 128 * 301 * 128 // 4)) * 4 for _ in calls]  # 16-byte aligned, anywhere in the first two thirds

    def run():
        got = []
        for (B, T, s1, s2, lam, masks), st in zip(calls, starts):
            out = arena[st:st + B * T * 128].view(B, T, 128)
            fe.forward_packed(pk, s1, s2, lam, masks, out=out, target_length=T)
            got.append(out.clone())
        torch.cuda.synchronize()
        return got

    fe.set_launch_overlap("off")
    want = run()
    for mode in ("stream", "call"):
        fe.set_launch_overlap(mode)
        for rep in range(3):
            got = run()
            for i, (g_, w_) in enumerate(zip(got, want)):
                assert torch.equal(g_, w_), (mode, rep, i, calls[i][0], calls[i][1])
    fe.close()


def test_c2_audioset_train_batch(fe):
    """C2: 100 clips, 150 waveforms, every even clip mixed, masks drawn like the reference (48 / 192)."""
    B = 100
    waves = [synth_wave(s, 160000, ("gauss", "am", "int16")[s % 3]) for s in range(150)]
    src1 = list(range(B))
    src2 = [100 + b // 2 if b % 2 == 0 else -1 for b in range(B)]
    lam_draws = np.random.default_rng(0).beta(10, 10, 50)
    lam = [float(lam_draws[b // 2]) if b % 2 == 0 else 0.0 for b in range(B)]
    torch.manual_seed(0)
    masks = [O.draw_mask_offsets(48, 192, 128, 1056) for _ in range(B)]
    out = fe(waves, src1, src2, lam, masks).cpu()
    torch.set_num_threads(max(1, min(16, os.cpu_count() or 1)))
    ref = O.batch_torch(waves, src1, src2, lam, masks, 1056, MEAN, STD)
    ref64 = O.batch_torch(waves, src1, src2, lam, masks, 1056, MEAN, STD, dtype=torch.float64)
    fill = _fill()
    dead = _dead_cells([160000] * B, masks, 1056)  # bit-exact mask / pad placement
    assert (out[dead] == fill).all() and (ref[dead] == fill).all()
    _same_fill_cells(out, ref, fill)
    err = (out - ref).abs() * STD
    ours64 = (out.double() - ref64).abs() * STD
    ref_err = (ref.double() - ref64).abs() * STD  # the fp32 reference's own distance to fp64 truth
    logmel64 = ref64 * STD + MEAN
    # Cells whose energy is within a factor ~20 of the float-epsilon floor (log-mel <= -13; typical cells are
    # -6 .. +3, the floor is -15.94) carry fp32 FFT round-off of relative size ~1e-3 in ANY fp32 implementation:
    # measured on this batch, the reference itself is > 1e-3 from fp64 truth on 4 such cells (max 1.5e-3).
    # Cells AT the floor (mel row 3 has no FFT bin at 16 kHz / 128 bins, so it is log(eps) in every live frame) are
    # not "near" it: they must come out bit-identical to the reference.
    at_floor = (ref == ((torch.log(torch.tensor(O.EPS32, dtype=torch.float32)) - MEAN) / STD)) & (ref != fill)
    assert at_floor.sum().item() >= 0.9 * (ref[:, :, 3] != fill).sum().item()
    assert torch.equal(out[at_floor], ref[at_floor])
    near_floor = (logmel64 <= -13.0) & (ref != fill) & ~at_floor
    assert err[~near_floor].max().item() <= TOL
    assert near_floor.sum().item() <= 2e-3 * err.numel()
    assert (err > TOL).sum().item() <= 2e-6 * err.numel()       # observed: 10 of 13.5 M
    assert err[near_floor].max().item() <= 4e-3
    # against fp64 truth we are as accurate as the fp32 reference is
    assert ours64.max().item() <= max(1.5 * ref_err.max().item(), TOL)
    assert ours64.pow(2).mean().sqrt().item() <= 1.25 * ref_err.pow(2).mean().sqrt().item()


def test_size_independent_properties_at_full_size(fe):
    """Properties that need no oracle: frame locality (prefix invariance), shift by whole frames, and
    invariance to a DC offset (removed per clip and per frame)."""
    x = synth_wave(80, 160000)
    full = fe([x]).cpu()
    pre = fe([x[:, :80400]]).cpu()  # 500 frames
    assert ((full[0, :500] - pre[0, :500]).abs() * STD).max() <= TOL  # clip mean differs -> rounding only
    sh = fe([x[:, 1600:]]).cpu()  # 10 frames later
    assert ((full[0, 10:998] - sh[0, :988]).abs() * STD).max() <= TOL
    # x + 0.25 re-quantises the input (ulp 3e-8 instead of <= 7.5e-9): two evaluations of different fp32 inputs, each within
    # TOL of the exact features of its own input => their difference is bounded by 2 TOL (observed: 0.6e-3 .. 1.0e-3 on the
    # weakest cell of the clip, depending on the rounding order of the kernel version), and by TOL for all but a few cells
    dc = fe([x + 0.25]).cpu()
    d = (full - dc).abs() * STD
    assert d.max() <= 2 * TOL
    assert (d > TOL).sum().item() <= 3


def test_dataset_drop_in_matches_reference_getitem(FE, golden_dir):
    """Same seeds -> our AudiosetDataset + GPU frontend reproduces the unmodified reference's __getitem__ output."""
    from psla_b200.dataloaders import AudiosetDataset, FrontendLoader
    for name in ("train", "eval", "skipnorm"):
        g = np.load(os.path.join(golden_dir, f"getitem_{name}.npz"))
        conf = json.loads(str(g["conf"]))
        waves = tiny_dataset_waves()
        store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
        with tempfile.TemporaryDirectory() as tmp:
            jpath, cpath = os.path.join(tmp, "t.json"), os.path.join(tmp, "l.csv")
            json.dump({"data": [{"wav": f"mem://{i}.wav", "labels": TINY_LABELS[i]} for i in range(len(waves))]}, open(jpath, "w"))
            with open(cpath, "w") as f:
                f.write("index,mid,display_name\n" + "".join(f"{i},{m},c{i}\n" for i, m in enumerate(TINY_CLASSES)))
            ds = AudiosetDataset(jpath, conf, cpath, loader=lambda p: (store[p].clone(), 16000), verbose=False)
            fe = FE(conf)
            items = []
            for seed in g["seeds"]:
                random.seed(int(seed)); np.random.seed(int(seed)); torch.manual_seed(int(seed))
                items += [ds[i] for i in range(len(ds))]
            batch = ds.collate(items)
            feats = FrontendLoader(None, fe).features(batch).cpu()
            ref = torch.from_numpy(g["feats"])
            div = 1.0 if conf.get("skip_norm") else conf["std"]
            fill = 0.0 if conf.get("skip_norm") else _fill(conf["mean"], conf["std"])
            _check(feats, ref, div=div, fill=fill)
            assert torch.equal(batch.labels, torch.from_numpy(g["labels"]))
            fe.close()


def test_c4_long_recording_chunks_equal_per_clip_oracle(FE):
    """C4: a long recording cut into 10 s windows; every window is an independent eval clip (own DC removal, no masks)."""
    from psla_b200.inference import LongRecordingFrontend
    fe = FE(CONF)
    n = 6 * 160000 + 12345
    x = synth_wave(90, n)[0] + 0.02  # DC offset: removed per window, not per recording
    lr = LongRecordingFrontend(fe, x, window_s=10.0)
    assert len(lr) == 7 and lr.lengths[-1] == 12345
    got = torch.cat([f for _, f in lr.batches(3)], dim=0)
    chunks = [x[s:s + l].unsqueeze(0) for s, l in zip(lr.starts.tolist(), lr.lengths.tolist())]
    ref = O.batch_torch(chunks, list(range(7)), [-1] * 7, [0.0] * 7, [(0, 0, 0, 0)] * 7, 1056, MEAN, STD)
    _check(got, ref, fill=_fill(), near_floor_below=-13.0)  # mel bin 0 (one FFT bin, tiny weight) dips to -14.9 in a few frames
    # batched == one window at a time, bit for bit; overlapping hop reuses the same device buffer
    one = torch.cat([lr.features(i, i + 1) for i in range(7)], dim=0)
    assert torch.equal(one, got)
    lr2 = LongRecordingFrontend(fe, x, window_s=10.0, hop_s=2.5)
    f2 = lr2.features(0, len(lr2))
    assert torch.equal(f2[0], got[0]) and torch.equal(f2[4], got[1])  # hop 2.5 s: window 4 starts at 10 s
    pred = lr.predict(lambda f: f.mean(dim=(1, 2)).unsqueeze(1), batch=4, aggregate=None)
    assert pred.per_window.shape == (7, 1) and pred.starts.tolist() == lr.starts.tolist() and pred.clip is None
    fe.close()


def test_c4_hour_long_recording_end_to_end(FE):
    """C4 at its real size: one hour of 16 kHz audio = 360 windows of 10 s, features against the oracle window by window,
    then the whole one-click path -- batched model inference (a stand-in for EffNetAttention: [b, T, 128] -> sigmoid
    [b, C]) one launch ahead on a second stream, per-window outputs, recording-level mean / max."""
    from psla_b200.inference import LongRecordingFrontend
    fe = FE(CONF)
    g = torch.Generator().manual_seed(7)
    n = 360 * 160000
    x = 0.1 * torch.randn(n, generator=g)
    x[5 * 160000:6 * 160000] *= 0.01   # a quiet window
    x[100 * 160000:101 * 160000] += 0.3  # a window with a DC offset
    lr = LongRecordingFrontend(fe, x, window_s=10.0)
    assert len(lr) == 360
    feats = torch.cat([f for _, f in lr.batches(128)], dim=0)
    torch.set_num_threads(max(1, min(16, os.cpu_count() or 1)))
    chunks = [x[s:s + l].unsqueeze(0) for s, l in zip(lr.starts.tolist(), lr.lengths.tolist())]
    ref = O.batch_torch(chunks, list(range(360)), [-1] * 360, [0.0] * 360, [(0, 0, 0, 0)] * 360, 1056, MEAN, STD)
    _check(feats, ref, fill=_fill(), near_floor_below=-13.0)
    C = 10
    wgt = torch.randn(128, C, generator=g).to(fe.device) * 0.2

    def model(f):  # time-average -> linear -> sigmoid, returns (output, attention) like EffNetAttention (:146)
        return torch.sigmoid(f.mean(dim=1) @ wgt), None

    want = torch.sigmoid(feats.mean(dim=1) @ wgt)
    for batch in (360, 128, 50):
        pred = lr.predict(model, batch=batch, aggregate="mean")
        assert torch.allclose(pred.per_window, want, rtol=0, atol=1e-6) and pred.per_window.shape == (360, C)
        assert torch.allclose(pred.clip, want.mean(0), atol=1e-5)
    pmax = lr.predict(model, batch=128, aggregate="max")
    assert torch.equal(pmax.clip, pmax.per_window.max(0).values)
    top = pmax.top(3, names=[f"class{i}" for i in range(C)])
    assert len(top) == 3 and top[0][1] >= top[1][1] >= top[2][1] and top[0][0].startswith("class")
    tl = pred.timeline(0)
    assert len(tl) == 360 and tl[1][0] == 10.0
    # a short last window weighs less in the mean
    lr2 = LongRecordingFrontend(fe, x[:2 * 160000 + 8000], window_s=10.0)
    p2 = lr2.predict(model, batch=2, aggregate="mean")
    fr = torch.tensor([998.0, 998.0, 48.0], device=fe.device)
    assert torch.allclose(p2.clip, (p2.per_window * fr.unsqueeze(1)).sum(0) / fr.sum(), atol=1e-6)
    fe.close()


def test_noise_roll_augmentation_matches_reference_getitem(FE, golden_dir):
    """audioset_dataset.py:197-199 through the drop-in dataset + loader: with the host-drawn U[0,1) fields the output
    equals the unmodified reference's __getitem__ (same seeds); with device-drawn fields only the distribution can."""
    from psla_b200.dataloaders import AudiosetDataset, FrontendLoader
    g = np.load(os.path.join(golden_dir, "getitem_noise.npz"))
    conf = json.loads(str(g["conf"]))
    waves = tiny_dataset_waves()
    store = {f"mem://{i}.wav": w for i, w in enumerate(waves)}
    with tempfile.TemporaryDirectory() as tmp:
        jpath, cpath = os.path.join(tmp, "t.json"), os.path.join(tmp, "l.csv")
        json.dump({"data": [{"wav": f"mem://{i}.wav", "labels": TINY_LABELS[i]} for i in range(len(waves))]}, open(jpath, "w"))
        with open(cpath, "w") as f:
            f.write("index,mid,display_name\n" + "".join(f"{i},{m},c{i}\n" for i, m in enumerate(TINY_CLASSES)))
        load = lambda p: (store[p].clone(), 16000)  # noqa: E731
        fe = FE(conf)
        outs = {}
        for mode in ("host", "device"):
            ds = AudiosetDataset(jpath, {**conf, "noise_field": mode}, cpath, loader=load, verbose=False)
            items = []
            for seed in g["seeds"]:
                random.seed(int(seed)); np.random.seed(int(seed)); torch.manual_seed(int(seed))
                items += [ds[i] for i in range(len(ds))]
            batch = ds.collate(items)
            outs[mode] = (FrontendLoader(None, fe).features(batch).cpu(), batch)
        feats, batch = outs["host"]
        ref = torch.from_numpy(g["feats"])
        assert ((feats - ref).abs() * conf["std"]).max().item() <= TOL
        assert torch.equal(batch.labels, torch.from_numpy(g["labels"]))
        # the kernel alone against torch ops on the same inputs: bit-exact (same two roundings per cell, exact row moves)
        base = fe.forward_packed(PackedWavesOf(batch, fe.device), batch.src1, batch.src2, batch.lam, batch.masks)
        want = torch.stack([torch.roll(base[b].cpu() + batch.noise[b] * float(np.float32(batch.noise_scale[b])) / 10, int(batch.roll[b]), 0)
                            for b in range(len(batch))])
        assert torch.equal(feats, want)
        # device-drawn fields: undo the roll, subtract the clean features -> U[0, scale / 10) per clip
        dfeats, dbatch = outs["device"]
        assert np.array_equal(dbatch.roll, batch.roll) and dbatch.noise is None
        for b in range(len(dbatch)):
            u = (torch.roll(dfeats[b], -int(dbatch.roll[b]), 0) - base[b].cpu()) * 10 / max(float(dbatch.noise_scale[b]), 1e-3)
            if dbatch.noise_scale[b] > 0.05:
                assert -1e-3 <= u.min().item() and u.max().item() <= 1.0 + 1e-3
                assert abs(u.mean().item() - 0.5) < 0.02 and abs(u.std().item() - 0.2887) < 0.02
        assert not torch.equal(dfeats[0], dfeats[len(waves)])  # different clips, different fields
        # argument checks of the C entry point
        with pytest.raises(Exception):
            fe.noise_roll(base, batch.noise_scale, batch.roll, out=base)  # in place
        with pytest.raises(Exception):
            fe.noise_roll(base, batch.noise_scale, np.full(len(batch), 96, np.int32))  # |shift| >= T
        fe.close()


def PackedWavesOf(batch, device):
    from psla_b200.frontend import PackedWaves
    return PackedWaves(batch.waves.wav.to(device), batch.waves.offsets, batch.waves.lengths)


@pytest.mark.parametrize("nmel", [64, 80, 40])
def test_other_num_mel_bins(FE, nmel):
    """`num_mel_bins` is an argument of the reference's dataset (audioset_dataset.py:63): sizes other than 128 take the
    generic (run-time tap count) mel path, rows of `nmel` floats, and the scalar path of the noise / roll kernel."""
    conf = {**CONF, "num_mel_bins": nmel, "target_length": 300}
    fe = FE(conf)
    waves = [synth_wave(700 + i, n) for i, n in enumerate((48000, 30000, 52345))]
    src1, src2, lam = [0, 1, 2, 0], [-1, 0, -1, 2], [0.0, 0.45, 0.0, 0.7]
    masks = [(0, 0, 0, 0), (5, min(20, nmel - 5), 40, 60), (nmel - 1, 1, 299, 1), (0, nmel // 2, 0, 0)]
    out = fe(waves, src1, src2, lam, masks)
    assert out.shape == (4, 300, nmel)
    ref = O.batch_torch(waves, src1, src2, lam, masks, 300, MEAN, STD, num_mel_bins=nmel)
    _check(out, ref, fill=_fill(), near_floor_below=-13.0)
    scale, roll = np.array([0.3, 0.0, 0.9, 0.5]), np.array([-10, 9, 0, 3], np.int32)
    field = torch.rand(4, 300, nmel)
    got = fe.noise_roll(out, scale, roll, field).cpu()
    want = torch.stack([torch.roll(out[b].cpu() + field[b] * float(np.float32(scale[b])) / 10, int(roll[b]), 0) for b in range(4)])
    assert torch.equal(got, want)
    fe.close()


@pytest.mark.parametrize("fast_log", [False, True])
@pytest.mark.parametrize("pcm16", [False, True])
def test_frequency_major_output_layouts(FE, fast_log, pcm16):
    """the consumer turns [B, T, F] into [B, 1, F, T] (src/models/Models.py:106-107) and autocast casts it
    to 16 bits; the "ft" / "ft_bf16" layouts write that directly.  Same arithmetic as the default layout: the fp32
    result is its transpose bit for bit, the bf16 one its round-to-nearest-even -- and the default layout is what the
    oracle checks.  Batches below and above the parameter-space descriptor limit (128 clips), masks, padding, mixup."""
    from psla_b200.frontend import pack_waveforms, to_pcm16, PackedWaves
    fe = FE(CONF, fast_log=fast_log)
    waves = [synth_wave(900 + i, n, kind) for i, (n, kind) in enumerate(
        ((160000, "gauss"), (160000, "am"), (100000, "int16"), (200000, "gauss"), (5000, "gauss"), (400, "gauss")))]
    packed = pack_waveforms(waves, device=fe.device)
    if pcm16:
        packed = PackedWaves(to_pcm16(packed.wav), packed.offsets, packed.lengths)
    for B in (7, 140):
        src1 = [b % 6 for b in range(B)]
        src2 = [{0: 1, 2: 3}.get(b % 6, -1) for b in range(B)]  # equal-length and longer (cut) partners: descriptors as kernel parameters
        if B > 128:
            src2[1] = 4  # a shorter partner (zero-padded tail) => clip statistics kernel + uploaded descriptors
        lam = [0.3 + 0.001 * b for b in range(B)]
        torch.manual_seed(B)
        masks = [O.draw_mask_offsets(48, 192, 128, 1056) for _ in range(B)]
        masks[0] = (0, 0, 0, 0)
        masks[2] = (0, 128, 0, 0)
        masks[3] = (100, 28, 1052, 4)
        fe.set_output_layout("tf")
        tf = fe.forward_packed(packed, src1, src2, lam, masks).cpu()
        fe.set_output_layout("ft")
        ft = fe.forward_packed(packed, src1, src2, lam, masks)
        assert ft.shape == (B, 128, 1056) and ft.dtype == torch.float32
        assert torch.equal(ft.cpu(), tf.transpose(1, 2))
        fe.set_output_layout("ft_bf16")
        fb = fe.forward_packed(packed, src1, src2, lam, masks)
        assert fb.shape == (B, 128, 1056) and fb.dtype == torch.bfloat16
        assert torch.equal(fb.cpu(), tf.transpose(1, 2).to(torch.bfloat16))
        if B == 7 and not pcm16:  # and directly against the oracle
            ref = O.batch_torch(waves, src1, src2, lam, masks, 1056, MEAN, STD)
            _check(ft.transpose(1, 2).contiguous(), ref, fill=_fill(), near_floor_below=-13.0)
            # host entry point: pinned in, pinned out, 16-bit elements
            host = PackedWaves(packed.wav.cpu().pin_memory(), packed.offsets, packed.lengths)
            out_host = torch.empty(B, 128, 1056, dtype=torch.bfloat16).pin_memory()
            fe.forward_host(host, src1, src2, lam, masks, out_host=out_host)
            torch.cuda.synchronize()
            assert torch.equal(out_host, fb.cpu())
            assert torch.equal(fe.forward_host(host, src1, src2, lam, masks).cpu(), fb.cpu())
    # what the layouts cannot do is an error at the call, not a wrong result
    with pytest.raises(Exception):
        fe.forward_packed(packed, [0], None, None, None, target_length=1054)
    with pytest.raises(Exception):
        fe.noise_roll(torch.zeros(1, 1056, 128, device=fe.device), [0.1], [1])
    fe.set_output_layout("tf")
    assert fe.forward_packed(packed, [0], None, None, None, target_length=1054).shape == (1, 1054, 128)
    fe.close()
    fe64 = FE({**CONF, "num_mel_bins": 64})
    fe64.set_output_layout("ft")
    with pytest.raises(Exception):
        fe64.forward_packed(packed, [0], None, None, None)
    fe64.close()


def test_batches_above_the_parameter_space_limit_go_out_in_chunks(FE):
    """More than 128 clips without a statistics pass: several launches of <= 128 clips (descriptors as kernel parameters,
    launches overlapping) instead of one with uploaded descriptors -- same bits as one launch and as clip-by-clip calls;
    a validation error anywhere in the batch is reported before anything is written."""
    fe = FE(CONF)
    waves = [synth_wave(1200 + i, n) for i, n in enumerate((160000, 120000, 160000, 52000, 16000))]
    from psla_b200.frontend import pack_waveforms
    packed = pack_waveforms(waves, device=fe.device)
    B = 300  # 3 chunks of 100
    src1 = [b % 5 for b in range(B)]
    src2 = [{1: 0, 3: 2}.get(b % 5, -1) for b in range(B)]  # partners at least as long: no statistics
    lam = [0.2 + 0.002 * b for b in range(B)]
    torch.manual_seed(3)
    masks = [O.draw_mask_offsets(48, 192, 128, 1056) for _ in range(B)]
    n0 = fe.launch_count
    chunked = fe.forward_packed(packed, src1, src2, lam, masks)
    assert fe.launch_count - n0 == 3
    fe.set_param_meta(False)  # one launch, descriptors uploaded
    n0 = fe.launch_count
    single = fe.forward_packed(packed, src1, src2, lam, masks)
    assert fe.launch_count - n0 == 1
    fe.set_param_meta(True)
    assert torch.equal(chunked, single)
    for b in (0, 99, 100, 101, 199, 200, 299):
        one = fe.forward_packed(packed, [src1[b]], [src2[b]], [lam[b]], [masks[b]])
        assert torch.equal(one[0], chunked[b])
    ref = O.batch_torch(waves, src1[:10], src2[:10], lam[:10], masks[:10], 1056, MEAN, STD)
    _check(chunked[:10], ref, fill=_fill())
    # an invalid clip in the LAST chunk: error, and no chunk has been launched
    out = torch.full((B, 1056, 128), 7.0, device=fe.device)
    bad = list(src1)
    bad[-1] = 99
    n0 = fe.launch_count
    with pytest.raises(Exception):
        fe.forward_packed(packed, bad, src2, lam, masks, out=out)
    torch.cuda.synchronize()
    assert fe.launch_count == n0 and bool((out == 7.0).all())
    # with a statistics clip in the batch (shorter partner) it stays one launch + the statistics kernel
    src2s = list(src2)
    src2s[7] = 4
    n0 = fe.launch_count
    stats = fe.forward_packed(packed, src1, src2s, lam, masks)
    assert fe.launch_count - n0 == 2
    keep = [b for b in range(B) if b != 7]
    assert torch.equal(stats[keep], chunked[keep])
    fe.close()
