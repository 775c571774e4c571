import os, sys, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from psla_b200 import params as P
from psla_b200.frontend import GpuFbankFrontend, PackedWaves
dev = torch.device('cuda', 0)
fe = GpuFbankFrontend({"num_mel_bins": 128, "target_length": 3000, "mean": -4.6476, "std": 4.5699}, device=dev)
for B in (24, 96):
    rng = np.random.default_rng(1)
    lens = rng.integers(4800, 480_000, B).astype(np.int64)
    lens[0], lens[1], lens[2] = 480_000, 500_000, 130_000
    starts = np.concatenate([[0], np.cumsum((lens + 3) // 4 * 4)[:-1]]).astype(np.int64)
    total = int(starts[-1] + lens[-1] + 8)
    sets = [PackedWaves(0.1 * torch.randn(total, device=dev), starts, lens) for _ in range(3)]
    outs = [torch.empty(B, 3000, 128, device=dev) for _ in range(3)]
    src1 = np.arange(B, dtype=np.int32)
    src2 = np.where(np.arange(B) % 2 == 0, (np.arange(B) * 7 + 3) % B, -1).astype(np.int32)
    lam = np.where(src2 >= 0, rng.beta(10, 10, B), 0.0)
    torch.manual_seed(3)
    masks = np.array([P.draw_masks(48, 192, 128, 3000) for _ in range(B)], dtype=np.int32)
    def step(k): fe.forward_packed(sets[k % 3], src1, src2, lam, masks, out=outs[k % 3])
    for mode in ("off", "call", "stream"):
        fe.set_launch_overlap(mode)
        for k in range(5): step(k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter(); e0.record()
        for k in range(100): step(k)
        e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
        print(B, mode, "gpu us/step %.1f" % (e0.elapsed_time(e1) * 10), "host us/call %.1f" % ((t1 - t0) * 1e4), flush=True)
    fe.set_timing(True)
    for k in range(30): step(k)
    a, b, n = fe.get_timing(); fe.set_timing(False)
    print(B, "serialised: stats kernel %.1f us, stream kernel %.1f us" % (a / n * 1e3, b / n * 1e3), "need stats:", int(((src2 >= 0) & (lens[np.maximum(src2, 0)] < lens)).sum()), flush=True)
    # without mixup (no stats pass)
    for k in range(5): fe.forward_packed(sets[k % 3], src1, None, None, masks, out=outs[k % 3])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(100): fe.forward_packed(sets[k % 3], src1, None, None, masks, out=outs[k % 3])
    e1.record(); torch.cuda.synchronize()
    print(B, "no mixup: gpu us/step %.1f" % (e0.elapsed_time(e1) * 10), flush=True)
