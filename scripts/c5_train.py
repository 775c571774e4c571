#!/usr/bin/env python
"""C5 (BASELINE.json config 5): the frontend feeding a training step of the reference's model family, end to end.

The model is OUT of scope of this repo (stock PyTorch) and only serves as the consumer: a
torchvision EfficientNet-B2 trunk with a 1-channel stem plus 4-head attention pooling over 527 classes -- the shape of
`EffNetAttention(b=2, head_num=4)` (src/models/Models.py:60-118; `efficientnet_pytorch` itself is not in the image) --
BCE loss, Adam, bf16 autocast, batch 100 per GPU, input `[B, 1056, 128]` -> `unsqueeze(1).transpose(2, 3)` (:106-107).

Arms (samples/s, 1 process per GPU; under torchrun every rank runs its own slice, gradients all-reduced by DDP):
  model_only       features already resident (the ceiling set by the model)
  gpu_frontend     pinned host 16-bit PCM batch -> psla_frontend_forward_host_pcm16 (features stay on the GPU) -> train step
  loader_frontend  16-bit wav files on tmpfs -> DataLoader workers (cores / ranks - 1 each) -> pinned shared slots ->
                   side-stream H2D -> fused kernel -> train step with the loader's own soft labels: the reference's loop
                   (src/traintest.py:81-110) with the dataloaders package swapped
  cpu_loader       the reference path: oracle port of AudiosetDataset.__getitem__ in ONE DataLoader with one worker per host
                   core -> .to(device) (src/traintest.py:81-85) -> train step.  At N > 1 the reference scales with
                   nn.DataParallel in a single process (src/traintest.py:44-45, egs/audioset/run.sh:6,68): rank 0 runs exactly
                   that over all N GPUs with a global batch of N x 100 while the other ranks wait.
Writes gpurun_out/c5.json (c5_n<N>.json under torchrun).
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
    st = {}

    def fresh_model():
        """Every arm trains its own freshly initialised model (same seed): the arms are independent short runs, and a
        randomly initialised network fed noise does not stay finite for hundreds of Adam steps."""
        torch.manual_seed(1)
        m = EffNetB2Attention().to(dev).to(memory_format=torch.channels_last)
        st["core"] = m
        st["model"] = nn.parallel.DistributedDataParallel(m, device_ids=[local]) if world > 1 else m
        # the reference warms the learning rate up from 0 over its first 1000 steps (traintest.py:90-95): these first few
        # dozen steps run at ~1e-6 (at 1e-4 from step 0 the randomly initialised stand-in goes non-finite within ten steps
        # once the labels change from step to step)
        st["opt"] = torch.optim.Adam(st["model"].parameters(), lr=1e-6, weight_decay=5e-7, betas=(0.95, 0.999))

    fresh_model()
    labels = (torch.rand(B, NCLS, device=dev) < 0.01).float()

    def train_step(feats, lab=None):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = st["model"](feats)
        loss = F.binary_cross_entropy(out.float().clamp(1e-7, 1 - 1e-7), labels if lab is None else lab)
        st["opt"].zero_grad(set_to_none=True)
        loss.backward()
        st["opt"].step()
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

    fresh_model()
    res["gpu_frontend"], res["gpu_frontend_loss"] = timed(gpu_step, steps)

    # ---- wav files -> DataLoader workers -> GPU frontend (the drop-in dataloaders package) ----
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import loader_bench as LB
    from psla_b200.dataloaders import AudiosetDataset, FrontendLoader, pcm16_lazy_loader
    cores = len(os.sched_getaffinity(0))
    workers = max(1, cores // world - 1)
    conf = {"num_mel_bins": NMEL, "target_length": T, "freqm": 48, "timem": 192, "mixup": 0.5, "dataset": "audioset",
            "mode": "train", "mean": MEAN, "std": STD, "noise": False}
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
    with tempfile.TemporaryDirectory(dir=base, prefix=f"psla_c5_{rank}_") as d:
        jpath, cpath = LB.make_files(d, 200, seed=rank)
        ds = AudiosetDataset(jpath, conf, label_csv=cpath, loader=pcm16_lazy_loader, verbose=False)
        sampler = torch.utils.data.WeightedRandomSampler(np.ones(len(ds)), 10 ** 7, replacement=True)  # run.py:84-85 (--bal)
        fl = FrontendLoader.create(ds, fe, batch_size=B, num_workers=workers, sampler=sampler, prefetch_factor=2)
        it = iter(fl)

        def loader_step():
            fb, lab = next(it)
            if os.environ.get("C5_DEBUG"):
                torch.cuda.synchronize()
                print("loader batch: features finite", bool(torch.isfinite(fb).all()), "range", float(fb.min()), float(fb.max()),
                      "labels range", float(lab.min()), float(lab.max()), "shape", tuple(lab.shape), flush=True)
                with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
                    o = st["model"](fb)
                print("  model output finite", bool(torch.isfinite(o).all()), float(o.float().min()), float(o.float().max()), flush=True)
            mode = os.environ.get("C5_LOADER_MODE", "")
            if mode == "ignore":
                return train_step(feats)
            if mode == "fixed_labels":
                return train_step(fb)
            if mode == "fixed_feats":
                return train_step(feats, lab)
            if mode == "clone_labels":
                return train_step(fb, lab.clone())
            return train_step(fb, lab)

        fresh_model()
        res["loader_frontend"], res["loader_frontend_loss"] = timed(loader_step, steps, warm=max(3, workers // 2))
        res["loader_workers_per_rank"] = workers
        res["loader_data_time_per_sample_us"] = fl.per_sample_data_time.avg * 1e6
        it.close()
        fl.close()
        del it, fl
    fe.close()

    # ---- the reference's CPU dataloader path: one process, one DataLoader over all host cores; nn.DataParallel at N > 1 ----
    if not os.environ.get("C5_NO_CPU"):
        barrier()
        if rank == 0:
            from oracle import cpu_baseline as CB
            workers = CB.host_cores()
            GB = world * B  # the reference's global batch: DataParallel splits it over the GPUs (traintest.py:44-45)
            it, sub = CB._loader(dict(CB.TRAIN_CONF), GB, workers, 40 * (GB + sub_pad(GB, workers)))
            if world > 1:
                torch.manual_seed(1)
                core = EffNetB2Attention().to(dev).to(memory_format=torch.channels_last)
                dp = nn.DataParallel(core, device_ids=list(range(world)))
                dp_opt = torch.optim.Adam(core.parameters(), lr=1e-6, weight_decay=5e-7, betas=(0.95, 0.999))
                dp_labels = (torch.rand(GB, NCLS, device=dev) < 0.01).float()
            else:
                fresh_model()
            cpu_steps = min(steps, 12 if world == 1 else 6)

            def cpu_step():
                parts, n = [], 0
                while n < GB:
                    fb, _ = next(it)
                    parts.append(fb)
                    n += fb.shape[0]
                f = torch.cat(parts)[:GB].to(dev, non_blocking=True)  # traintest.py:84
                if world == 1:
                    return train_step(f)
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    out = dp(f)
                loss = F.binary_cross_entropy(out.float().clamp(1e-7, 1 - 1e-7), dp_labels)
                dp_opt.zero_grad(set_to_none=True)
                loss.backward()
                dp_opt.step()
                return loss

            for _ in range(2):
                cpu_step()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(cpu_steps):
                loss = cpu_step()
            loss.item()
            for i in range(world):
                torch.cuda.synchronize(i)
            res["cpu_loader"] = GB * cpu_steps / (time.perf_counter() - t0)
            res["cpu_loader_workers"] = workers
            res["cpu_loader_mode"] = "single process" + (f", nn.DataParallel over {world} GPUs, global batch {GB}" if world > 1 else "")
        barrier()
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
