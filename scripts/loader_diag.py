#!/usr/bin/env python
"""Where the worker-fed path spends its time: pure DataLoader rate (slots released at once), H2D rate out of a pool slot,
and the full FrontendLoader loop with a per-batch breakdown.  python scripts/loader_diag.py [workers ...]"""
import functools, json, os, sys, tempfile, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import loader_bench as LB
from psla_b200.dataloaders import AudiosetDataset, FrontendLoader, PinnedSlotPool, pcm16_lazy_loader
from psla_b200.frontend import GpuFbankFrontend

conf = {"num_mel_bins": 128, "target_length": 1056, "freqm": 48, "timem": 192, "mixup": 0.5, "dataset": "audioset",
        "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
dev = torch.device("cuda", 0)
out = {}
with tempfile.TemporaryDirectory(dir="/dev/shm") as d:
    j, c = LB.make_files(d, 200)
    ds = AudiosetDataset(j, conf, label_csv=c, loader=pcm16_lazy_loader, verbose=False)
    for W in [int(a) for a in sys.argv[1:]] or [4, 8, 15]:
        pool = PinnedSlotPool(2 * W + 3, 2 * 100 * 160004, torch.int16)
        samp = torch.utils.data.RandomSampler(ds, replacement=True, num_samples=10 ** 9)
        dl = torch.utils.data.DataLoader(ds, batch_size=100, num_workers=W, sampler=samp, prefetch_factor=2,
                                         collate_fn=functools.partial(AudiosetDataset.collate, pool=pool))
        it = iter(dl)
        for _ in range(W):
            b = next(it); pool.release(b.slot)
        t0 = time.perf_counter(); n = 0
        while time.perf_counter() - t0 < 2.5:
            b = next(it); pool.release(b.slot); n += len(b)
        rate = n / (time.perf_counter() - t0)
        # H2D out of a slot
        v = pool.view(0, 100 * 240000)
        dst = torch.empty_like(v, device=dev)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        for _ in range(10):
            dst.copy_(v, non_blocking=True)
        torch.cuda.synchronize()
        gbs = 10 * v.numel() * 2 / (time.perf_counter() - t1) / 1e9
        out[f"w{W}"] = {"dataloader_only_clips_s": round(rate), "slot_is_pinned": bool(v.is_pinned()), "registered": pool.registered,
                        "h2d_GBs": round(gbs, 1)}
        pool.close(); del it, dl
        print(json.dumps(out[f"w{W}"]), flush=True)
    # full loop with a breakdown
    W = 8
    fe = GpuFbankFrontend(conf, device=dev)
    samp = torch.utils.data.RandomSampler(ds, replacement=True, num_samples=10 ** 9)
    fl = FrontendLoader.create(ds, fe, batch_size=100, num_workers=W, sampler=samp, prefetch_factor=2)
    it = iter(fl.loader)
    t_next = t_feat = t_reap = 0.0; n = 0
    for k in range(W + 40):
        a = time.perf_counter(); fl._reap(block=True); b0 = time.perf_counter()
        batch = next(it); b1 = time.perf_counter()
        fb, lab = fl.features(batch, with_labels=True); b2 = time.perf_counter()
        if k >= W:
            t_reap += b0 - a; t_next += b1 - b0; t_feat += b2 - b1; n += 1
    torch.cuda.synchronize()
    print(json.dumps({"full_w8_ms_per_batch": {"reap": t_reap / n * 1e3, "next": t_next / n * 1e3, "features": t_feat / n * 1e3}}), flush=True)
    # statement-level timing of one staged batch (same calls as FrontendLoader._stage / features)
    import numpy as np
    from psla_b200.frontend import PackedWaves
    T = {}
    def tick(name, t0):
        T[name] = T.get(name, 0.0) + time.perf_counter() - t0
    cs = torch.cuda.Stream(device=dev)
    devbuf = [torch.empty(64 << 20, dtype=torch.uint8, device=dev) for _ in range(3)]
    ring = torch.empty(100 * 527, dtype=torch.float32, pin_memory=True)
    n = 0
    for k in range(30):
        fl._reap(block=True)
        batch = next(it)
        t0 = time.perf_counter(); nn = int(batch.waves.offsets[-1] + (batch.waves.lengths[-1] + 3) // 4 * 4); wav = fl.pool.view(batch.slot, nn); tick("view", t0)
        t0 = time.perf_counter(); dst = devbuf[k % 3][:nn * 2].view(torch.int16); tick("dst", t0)
        t0 = time.perf_counter(); hl = fl.host_labels(batch); ring[:hl.numel()].copy_(hl.reshape(-1)); tick("labels_ring", t0)
        t0 = time.perf_counter()
        with torch.cuda.stream(cs):
            dst.copy_(wav, non_blocking=True)
            tick("h2d_call", t0)
            t0 = time.perf_counter(); lab = ring[:hl.numel()].to(dev, non_blocking=True); tick("labels_h2d", t0)
            t0 = time.perf_counter(); ev = torch.cuda.Event(); ev.record(cs); tick("event", t0)
        t0 = time.perf_counter(); torch.cuda.current_stream(dev).wait_event(ev); tick("wait_event", t0)
        t0 = time.perf_counter(); fb = fe.forward_packed(PackedWaves(dst, batch.waves.offsets, batch.waves.lengths), batch.src1, batch.src2, batch.lam, batch.masks); tick("forward_packed", t0)
        t0 = time.perf_counter(); ev.synchronize(); tick("copy_wait", t0)
        fl.pool.release(batch.slot)
        n += 1
    torch.cuda.synchronize()
    print(json.dumps({k: round(v / n * 1e3, 3) for k, v in T.items()}), flush=True)
    fl.close()
