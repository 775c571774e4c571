#!/usr/bin/env python
"""SURVEY 8f row 3 measured: the bench workload (AudioSet training batch, device-resident inputs, back-to-back launches)
with the three output layouts, next to what the consumer pays without the fusion -- the default [B, T, F] fp32 output
followed by the reference's `unsqueeze(1).transpose(2, 3)` made contiguous (what the first convolution needs) and the
autocast cast to bf16.  Writes gpurun_out/layout_bench.json."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

B, T, NMEL, N, NWAV, NSETS = 100, 1056, 128, 160000, 150, 4


def main():
    from psla_b200 import params as P
    from psla_b200.frontend import GpuFbankFrontend, PackedWaves
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device=dev).manual_seed(1234)
    offs, lens = np.arange(NWAV, dtype=np.int64) * N, np.full(NWAV, N, dtype=np.int64)
    sets = [PackedWaves(0.1 * torch.randn(NWAV * N, device=dev, generator=gen), offs, lens) for _ in range(NSETS)]
    b = np.arange(B)
    src1, src2 = b.astype(np.int32), np.where(b % 2 == 0, B + b // 2, -1).astype(np.int32)
    lam = np.where(b % 2 == 0, np.random.default_rng(0).beta(10, 10, B), 0.0)
    torch.manual_seed(0)
    masks = np.array([P.draw_masks(48, 192, NMEL, T) for _ in range(B)], dtype=np.int32)
    fe = GpuFbankFrontend({"num_mel_bins": NMEL, "target_length": T, "mean": -4.6476, "std": 4.5699}, device=dev)
    fe.set_launch_overlap("stream")  # back-to-back calls on one stream with nothing in between (like bench.py's value leg)

    def timed(step, steps=200):
        for k in range(10):
            step(k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(steps):
            step(k)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / steps

    res = {}
    for layout in ("tf", "ft", "ft_bf16"):
        fe.set_output_layout(layout)
        shape, dtype = fe.out_shape_dtype(B, T)
        outs = [torch.empty(shape, dtype=dtype, device=dev) for _ in range(NSETS)]
        step = lambda k: fe.forward_packed(sets[k % NSETS], src1, src2, lam, masks, out=outs[k % NSETS])
        ms = min(timed(step) for _ in range(3))
        res[layout] = {"ms_per_launch": ms, "clips_per_s": B / (ms * 1e-3), "out_bytes": outs[0].numel() * outs[0].element_size()}
        if layout == "tf":
            # the consumer's side of the unfused path: [B, T, F] -> contiguous [B, 1, F, T] (-> bf16)
            def unfused(k, cast):
                o = fe.forward_packed(sets[k % NSETS], src1, src2, lam, masks, out=outs[k % NSETS])
                x = o.unsqueeze(1).transpose(2, 3).contiguous()
                return x.to(torch.bfloat16) if cast else x
            ms_t = min(timed(lambda k: unfused(k, False)) for _ in range(3))
            ms_c = min(timed(lambda k: unfused(k, True)) for _ in range(3))
            res["tf_then_torch_transpose"] = {"ms_per_step": ms_t, "clips_per_s": B / (ms_t * 1e-3)}
            res["tf_then_torch_transpose_bf16"] = {"ms_per_step": ms_c, "clips_per_s": B / (ms_c * 1e-3)}
        print(layout, res[layout], flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "layout_bench.json"), "w"), indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
