#!/usr/bin/env python
"""C5 (BASELINE.json config 5): the frontend feeding a training step of the reference's model family, end to end.

The model is OUT of scope of this repo (stock PyTorch) and only serves as the consumer: a
torchvision EfficientNet-B2 trunk with a 1-channel stem plus 4-head attention pooling over 527 classes -- the shape of
`EffNetAttention(b=2, head_num=4)` (src/models/Models.py:60-118; `efficientnet_pytorch` itself is not in the image) --
BCE loss, Adam, bf16 autocast, batch 100 per GPU, input `[B, 1056, 128]` -> `unsqueeze(1).transpose(2, 3)` (:106-107).

Arms (samples/s, 1 process per GPU; under torchrun every rank runs its own slice, gradients all-reduced by DDP):
  gpu_frontend   pinned host 16-bit PCM batch -> psla_frontend_forward_host_pcm16 (features stay on the GPU) -> train step
  model_only     features already resident (the ceiling set by the model)
  cpu_loader     the reference path: oracle port of AudiosetDataset.__getitem__ in a DataLoader with one worker per
                 host core -> .to(device) (src/traintest.py:81-85) -> train step           (rank 0 / N = 1 only)
Writes gpurun_out/c5.json.
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

B, T, NMEL, NCLS, N = 100, 1056, 128, 527, 160000
MEAN, STD = -4.6476, 4.5699


class MultiHeadAttentionPool(nn.Module):
    """Same computation as the reference's MHeadAttention (src/models/HigherModels.py): per head a sigmoid attention map
    and a sigmoid classification map over the time axis, attention normalised over time, heads combined by learnt weights."""

    def __init__(self, n_in, n_out, heads=4):
        super().__init__()
        self.att = nn.ModuleList([nn.Conv2d(n_in, n_out, 1) for _ in range(heads)])
        self.cla = nn.ModuleList([nn.Conv2d(n_in, n_out, 1) for _ in range(heads)])
        self.head_weight = nn.Parameter(torch.full((heads,), 1.0 / heads))

    def forward(self, x):  # [B, C, 1, T']
        outs = []
        for att, cla, w in zip(self.att, self.cla, self.head_weight):
            a = torch.sigmoid(att(x)).clamp(1e-7, 1 - 1e-7)
            a = a / a.sum(dim=3, keepdim=True)
            outs.append((a * torch.sigmoid(cla(x))).sum(dim=3)[:, :, 0] * w)
        return torch.stack(outs, 0).sum(0)


class EffNetB2Attention(nn.Module):
    def __init__(self):
        super().__init__()
        import torchvision
        trunk = torchvision.models.efficientnet_b2(weights=None).features
        stem = trunk[0][0]
        trunk[0][0] = nn.Conv2d(1, stem.out_channels, stem.kernel_size, stem.stride, stem.padding, bias=False)
        self.trunk = trunk
        self.pool = MultiHeadAttentionPool(1408, NCLS, 4)
        self.freq_pool = nn.AvgPool2d((4, 1))

    def forward(self, x):  # [B, T, 128]
        x = x.unsqueeze(1).transpose(2, 3)  # [B, 1, 128, T]  (Models.py:106-107)
        x = self.trunk(x)                   # [B, 1408, 4, T/32]
        x = self.freq_pool(x)               # [B, 1408, 1, T/32]
        return self.pool(x)


def main():
    import torch.distributed as dist
    from psla_b200 import params as P
    from psla_b200.frontend import GpuFbankFrontend, PackedWaves
    world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1 + rank)
    model = EffNetB2Attention().to(dev).to(memory_format=torch.channels_last)
    if world > 1:
        model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
    opt = torch.optim.Adam(model.parameters(), lr=1e-4, weight_decay=5e-7, betas=(0.95, 0.999))
    labels = (torch.rand(B, NCLS, device=dev) < 0.01).float()

    def train_step(feats):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(feats)
        loss = F.binary_cross_entropy(out.float().clamp(1e-7, 1 - 1e-7), labels)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warm=3):
        for _ in range(warm):
            step_fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            loss = step_fn()
        loss.item()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        return world * B * steps / dt, float(loss.item())

    res = {"n_gpus": world, "batch_per_gpu": B, "model": "torchvision efficientnet_b2 trunk (1-ch stem) + 4-head attention pooling, 527 classes, "
           "bf16 autocast, Adam, BCE", "device": torch.cuda.get_device_name(local)}
    steps = int(os.environ.get("C5_STEPS", "20"))

    # ---- model only ----
    feats = torch.randn(B, T, NMEL, device=dev)
    res["model_only"], _ = timed(lambda: train_step(feats), steps)

    # ---- GPU frontend from pinned host PCM ----
    fe = GpuFbankFrontend({"num_mel_bins": NMEL, "target_length": T, "mean": MEAN, "std": STD}, device=dev)
    n_wav = 150
    offs, lens = np.arange(n_wav, dtype=np.int64) * N, np.full(n_wav, N, dtype=np.int64)
    g = torch.Generator().manual_seed(7 + rank)
    pcm_sets = [PackedWaves(torch.clamp(torch.round(0.1 * torch.randn(n_wav * N, generator=g) * 32768), -32768, 32767).to(torch.int16).pin_memory(),
                            offs, lens) for _ in range(3)]
    b = np.arange(B)
    src1, src2 = b.astype(np.int32), np.where(b % 2 == 0, B + b // 2, -1).astype(np.int32)
    lam = np.where(b % 2 == 0, np.random.default_rng(0).beta(10, 10, B), 0.0)
    state = {"k": 0}

    def gpu_step():
        torch.manual_seed(state["k"])
        masks = np.array([P.draw_masks(48, 192, NMEL, T) for _ in range(B)], dtype=np.int32)  # host draws, reference order
        f = fe.forward_host(pcm_sets[state["k"] % 3], src1, src2, lam, masks)
        state["k"] += 1
        return train_step(f)

    res["gpu_frontend"], res["gpu_frontend_loss"] = timed(gpu_step, steps)
    fe.close()

    # ---- the reference's CPU dataloader path (rank 0, N = 1) ----
    if world == 1 and not os.environ.get("C5_NO_CPU"):
        from oracle import cpu_baseline as CB
        workers = CB.host_cores()
        it, sub = CB._loader(dict(CB.TRAIN_CONF), B, workers, 40 * (B + sub_pad(B, workers)))

        def cpu_step():
            parts, n = [], 0
            while n < B:
                fb, _ = next(it)
                parts.append(fb)
                n += fb.shape[0]
            f = torch.cat(parts)[:B].to(dev, non_blocking=True)  # traintest.py:84
            return train_step(f)

        res["cpu_loader"], _ = timed(cpu_step, min(steps, 12))
        res["cpu_loader_workers"] = workers
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        name = "c5.json" if world == 1 else f"c5_n{world}.json"
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", name), "w"), indent=1)
        print(json.dumps(res))
    if world > 1:
        dist.destroy_process_group()


def sub_pad(batch, workers):
    return max(1, batch // max(workers, 1))


if __name__ == "__main__":
    main()
