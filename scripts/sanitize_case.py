#!/usr/bin/env python
"""Small workload for compute-sanitizer (scripts/gpu_sanitize.sh): the smoke batch plus a C2-shape (mixup + masks) and a
C3-shape (ragged lengths, shorter partners -> statistics kernel) batch, fp32 and int16 PCM input, all three output layouts,
both descriptor paths, exact clip means, overlapping launches, and the noise / roll kernel.  Sizes are small: the tools
slow kernels down by two orders of magnitude."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from psla_b200 import params as P  # noqa: E402
from psla_b200.frontend import GpuFbankFrontend, pack_waveforms, to_pcm16  # noqa: E402
from psla_b200.synth import synth_wave  # noqa: E402

dev = torch.device("cuda", 0)
T = 200
conf = {"num_mel_bins": 128, "target_length": T, "mean": -4.6476, "std": 4.5699}
waves = [synth_wave(s, n, ("gauss", "am", "int16")[s % 3]) for s, n in enumerate((32000, 32000, 20000, 47000, 9000, 32000, 16000, 401))]
B = 8
src1 = np.arange(B, dtype=np.int32)
src2 = np.array([1, -1, 3, 2, -1, 4, 0, -1], dtype=np.int32)  # longer and SHORTER partners (the latter need clip statistics)
lam = np.random.default_rng(0).beta(10, 10, B)
torch.manual_seed(0)
masks = np.array([P.draw_masks(48, 60, 128, T) for _ in range(B)], dtype=np.int32)
n_calls = 0
for exact in (False, True):
    fe = GpuFbankFrontend(conf, device=dev, exact_clip_mean=exact)
    for pcm in (False, True):
        pk = pack_waveforms([to_pcm16(w) for w in waves] if pcm else waves, device=dev)
        for layout in (("tf", "ft", "ft_bf16") if not exact else ("tf",)):
            fe.set_output_layout(layout)
            for pmeta in (True, False):
                fe.set_param_meta(pmeta)
                out = fe.forward_packed(pk, src1, src2, lam, masks)
                out2 = fe.forward_packed(pk, src1[:3], None, None, None)  # eval clips: no statistics pass
                n_calls += 2
        fe.set_output_layout("tf")
        fe.set_param_meta(True)
    # overlapping launches (PSLA_OVERLAP_STREAM): back-to-back calls into rotating buffers
    fe.set_launch_overlap("stream")
    pk = pack_waveforms(waves, device=dev)
    bufs = [torch.empty(3, T, 128, device=dev) for _ in range(3)]
    for k in range(9):
        fe.forward_packed(pk, src1[k % 5:k % 5 + 3], None, None, None, out=bufs[k % 3])
        n_calls += 1
    fe.set_launch_overlap("call")
    feats = fe.forward_packed(pk, src1, None, None, None)
    fe.noise_roll(feats, np.full(B, 0.5), np.arange(B) - 4, None, seed=1)
    fe.noise_roll(feats, np.full(B, 0.5), np.arange(B) - 4, torch.rand(B, T, 128), seed=1)
    torch.cuda.synchronize()
print(f"sanitize_case ok: {n_calls} forward calls")
