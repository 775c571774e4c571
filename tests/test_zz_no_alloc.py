"""psla_frontend_forward_static (include/psla_frontend.h): the allocation-free entry point, called through the C ABI and
checked against psla_frontend_forward on the same inputs.  `-m gpu`.  Collected last on purpose
(file name).

Both entry points launch the same kernels; the only difference is where the statistics kernel's partial sums live (the
caller's workspace vs the library's static device array), so the outputs must be bit-identical -- asserted as such, which
is inside the 1e-3 log-mel tolerance of tests/test_frontend_gpu.py.  Every batch here needs the statistics pass (mixup
partners shorter than their clip, or exact_clip_mean mode): that is the path that uses the scratch.
"""
import numpy as np
import pytest
import torch

from psla_b200.synth import synth_wave

pytestmark = pytest.mark.gpu

MEAN, STD, T = -4.6476, 4.5699, 1056
CONF = {"num_mel_bins": 128, "target_length": T, "mean": MEAN, "std": STD}
TOL = 1e-3
STATIC_MAX_CLIPS = 128  # PSLA_STATIC_MAX_CLIPS
E_STATIC_LIMIT = -8


@pytest.fixture(scope="module")
def FE():
    from psla_b200 import build
    build.build()
    from psla_b200.frontend import GpuFbankFrontend
    return GpuFbankFrontend


def _batch(seed, n_clips, dev):
    """Ragged clips, every second one mixed with a SHORTER partner (zero-padded tail -> whole-clip mean of the partner
    -> statistics pass), SpecAugment masks on some."""
    from psla_b200.frontend import pack_waveforms

    rng = np.random.default_rng(seed)
    lengths = [int(rng.integers(60_000, 160_001)) for _ in range(n_clips)]
    waves = [synth_wave(seed * 1000 + i, n, ("gauss", "am", "int16")[i % 3]) for i, n in enumerate(lengths)]
    partners = [synth_wave(seed * 1000 + 500 + i, max(401, lengths[2 * i] - int(rng.integers(1, 40_000))), "gauss") for i in range((n_clips + 1) // 2)]
    packed = pack_waveforms(waves + partners, device=dev)
    src1 = np.arange(n_clips, dtype=np.int32)
    src2 = np.array([n_clips + b // 2 if b % 2 == 0 else -1 for b in range(n_clips)], dtype=np.int32)
    lam = np.where(src2 >= 0, rng.beta(10, 10, n_clips), 0.0).astype(np.float64)
    masks = np.zeros((n_clips, 4), dtype=np.int32)
    for b in range(0, n_clips, 3):
        masks[b] = (int(rng.integers(0, 80)), int(rng.integers(0, 48)), int(rng.integers(0, 300)), int(rng.integers(0, 192)))
    return packed, src1, src2, lam, masks


def _call_static(fe, packed, src1, src2, lam, masks, out):
    """The raw C call: device pointers of `packed.wav` / `out`, host pointers of the numpy arrays, the current stream."""
    B = len(src1)
    off = np.ascontiguousarray(packed.offsets, dtype=np.int64)
    length = np.ascontiguousarray(packed.lengths, dtype=np.int64)
    keep = (off, length, src1, src2, lam, masks)  # alive until the call returns
    st = fe.lib.psla_frontend_forward_static(
        fe._h, packed.wav.data_ptr(), packed.wav.numel(), off.ctypes.data, length.ctypes.data, len(off), src1.ctypes.data,
        src2.ctypes.data, lam.ctypes.data, masks.ctypes.data, B, T, fe.norm_mean, fe.norm_div, int(fe.skip_norm),
        out.data_ptr(), torch.cuda.current_stream(fe.device).cuda_stream)
    del keep
    return st


@pytest.mark.parametrize("exact", [False, True], ids=["fused-mean", "exact-mean"])
def test_forward_static_matches_forward_on_calls_in_a_row(FE, exact):
    fe = FE(CONF, exact_clip_mean=exact)
    dev = fe.device
    batches = [_batch(11, 12, dev), _batch(12, 7, dev)]
    # reference results first: the existing entry point with a caller-provided workspace
    want = [fe.forward_packed(pk, s1, s2, lam, mk).clone() for pk, s1, s2, lam, mk in batches]
    torch.cuda.synchronize(dev)
    launches0 = fe.launch_count
    # five calls in a row on one stream, nothing between them (the three partial-sum regions of the static array are
    # reused from the fourth call on, with the stale sums of a DIFFERENT batch in them), outputs checked afterwards
    order = [0, 1, 0, 1, 1]
    outs = [torch.full((len(batches[k][1]), T, 128), float("nan"), device=dev) for k in order]
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    for k, out in zip(order, outs):
        assert _call_static(fe, *batches[k], out) == 0
    torch.cuda.synchronize(dev)
    free1 = torch.cuda.mem_get_info(dev)[0]
    assert fe.launch_count - launches0 == 2 * len(order)  # statistics kernel + streaming kernel per call: the scratch was used
    assert free1 >= free0, "device memory in use grew across forward_static calls"
    for k, out in zip(order, outs):
        assert not torch.isnan(out).any()
        assert ((out - want[k]).abs() * STD).max().item() <= TOL
        assert torch.equal(out, want[k])
    # and the two entry points interleaved on the same handle
    out = torch.empty_like(want[0])
    got_ws = fe.forward_packed(*batches[0])
    assert _call_static(fe, *batches[0], out) == 0
    torch.cuda.synchronize(dev)
    assert torch.equal(out, want[0]) and torch.equal(got_ws, want[0])


def test_forward_static_limits(FE):
    fe = FE(CONF)
    dev = fe.device
    n = STATIC_MAX_CLIPS + 1
    wav = synth_wave(5, 16000, "gauss")
    from psla_b200.frontend import pack_waveforms
    pk = pack_waveforms([wav], device=dev)
    src1 = np.zeros(n, dtype=np.int32)
    src2 = np.full(n, -1, dtype=np.int32)
    lam = np.zeros(n, dtype=np.float64)
    masks = np.zeros((n, 4), dtype=np.int32)
    out = torch.zeros(n, T, 128, device=dev)
    l0 = fe.launch_count
    assert _call_static(fe, pk, src1, src2, lam, masks, out) == E_STATIC_LIMIT  # one clip over the maximum: refused, nothing launched
    assert fe.launch_count == l0
    msg = fe.lib.psla_strerror(E_STATIC_LIMIT).decode()
    assert "128" in msg and "forward_static" in msg
    # exactly the maximum is served, and equals forward()
    m = STATIC_MAX_CLIPS
    assert _call_static(fe, pk, src1[:m].copy(), src2[:m].copy(), lam[:m].copy(), masks[:m].copy(), out[:m]) == 0
    want = fe.forward_packed(pk, src1[:m], None, None, None)
    torch.cuda.synchronize(dev)
    assert torch.equal(out[:m], want)
    # a handle mode whose launches would allocate (pinned descriptor slot) is refused as well
    fe.set_param_meta(False)
    try:
        assert _call_static(fe, pk, src1[:4].copy(), src2[:4].copy(), lam[:4].copy(), masks[:4].copy(), out[:4]) == E_STATIC_LIMIT
    finally:
        fe.set_param_meta(True)
