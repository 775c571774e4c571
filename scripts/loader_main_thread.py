#!/usr/bin/env python
"""Main-process time per batch of the worker-fed path at full worker count: waiting for / receiving the next batch, releasing
slots, staging + launching.  python scripts/loader_main_thread.py [workers]"""
import json, os, sys, tempfile, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import loader_bench as LB
from psla_b200.dataloaders import AudiosetDataset, FrontendLoader, pcm16_lazy_loader
from psla_b200.frontend import GpuFbankFrontend

W = int(sys.argv[1]) if len(sys.argv) > 1 else max(1, len(os.sched_getaffinity(0)) - 1)
conf = {"num_mel_bins": 128, "target_length": 1056, "freqm": 48, "timem": 192, "mixup": 0.5, "dataset": "audioset",
        "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
dev = torch.device("cuda", 0)
with tempfile.TemporaryDirectory(dir="/dev/shm") as d:
    j, c = LB.make_files(d, 200)
    ds = AudiosetDataset(j, conf, label_csv=c, loader=pcm16_lazy_loader, verbose=False)
    fe = GpuFbankFrontend(conf, device=dev)
    samp = torch.utils.data.RandomSampler(ds, replacement=True, num_samples=10 ** 9)
    fl = FrontendLoader.create(ds, fe, batch_size=100, num_workers=W, sampler=samp, prefetch_factor=2)
    it = iter(fl.loader)
    t_next = t_feat = t_reap = 0.0
    n = 0
    t_start = None
    for k in range(10 ** 9):
        a = time.perf_counter(); fl._reap(block=True, keep=2); b0 = time.perf_counter()
        batch = next(it); b1 = time.perf_counter()
        fb, lab = fl.features(batch, with_labels=True); b2 = time.perf_counter()
        if k == 2 * W:
            t_start = time.perf_counter()
        if k > 2 * W:
            t_reap += b0 - a; t_next += b1 - b0; t_feat += b2 - b1; n += 1
            if time.perf_counter() - t_start > 3.0:
                break
    torch.cuda.synchronize()
    dt = time.perf_counter() - t_start
    print(json.dumps({"workers": W, "clips_per_s": round(n * 100 / dt), "ms_per_batch": {"reap": round(t_reap / n * 1e3, 3), "next": round(t_next / n * 1e3, 3),
                                                                                          "features": round(t_feat / n * 1e3, 3), "total": round(dt / n * 1e3, 3)}}), flush=True)
    fl.close()
