import os, sys, tempfile, torch, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'scripts'))
import loader_bench as LB
from psla_b200 import params as P
from psla_b200.dataloaders import AudiosetDataset, FrontendLoader, pcm16_lazy_loader
from psla_b200.frontend import GpuFbankFrontend, PackedWaves
dev = torch.device('cuda', 0)
conf = {"num_mel_bins": 128, "target_length": 1056, "freqm": 48, "timem": 192, "mixup": 0.5, "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
fe = GpuFbankFrontend(conf, device=dev)
# 1) full-shape guard bands around out, many mask draws
B, T, N = 100, 1056, 160000
n_wav = 150
offs, lens = np.arange(n_wav, dtype=np.int64) * N, np.full(n_wav, N, dtype=np.int64)
wav = (0.1 * torch.randn(n_wav * N, device=dev))
pk = PackedWaves(wav, offs, lens)
b = np.arange(B); src1 = b.astype(np.int32); src2 = np.where(b % 2 == 0, B + b // 2, -1).astype(np.int32)
lam = np.where(b % 2 == 0, 0.4, 0.0)
n_out = B * T * 128
big = torch.full((n_out + 2 * (1 << 20),), -777.0, device=dev)
out = big[1 << 20:(1 << 20) + n_out].view(B, T, 128)
for it in range(30):
    torch.manual_seed(it)
    masks = np.array([P.draw_masks(48, 192, 128, T) for _ in range(B)], dtype=np.int32)
    fe.forward_packed(pk, src1, src2, lam, masks, out=out)
torch.cuda.synchronize()
print("guard bands intact:", bool((big[:1 << 20] == -777.0).all() and (big[(1 << 20) + n_out:] == -777.0).all()), "out finite", bool(torch.isfinite(out).all()), flush=True)
# 2) canaries in the caching allocator while the loader runs with a consumer that allocates and frees
canaries = [torch.full((1 << 18,), float(i), device=dev) for i in range(64)]
with tempfile.TemporaryDirectory(dir='/dev/shm') as d:
    j, c = LB.make_files(d, 200, seed=0)
    ds = AudiosetDataset(j, conf, label_csv=c, loader=pcm16_lazy_loader, verbose=False)
    sampler = torch.utils.data.WeightedRandomSampler(np.ones(len(ds)), 10 ** 7, replacement=True)
    fl = FrontendLoader.create(ds, fe, batch_size=100, num_workers=15, sampler=sampler, prefetch_factor=2)
    it = iter(fl)
    w = torch.randn(128, 527, device=dev, requires_grad=True)
    bad_at = None
    for k in range(60):
        fb, lab = next(it)
        tmp = [torch.empty(int(np.random.randint(1, 64)) << 18, device=dev).fill_(1.0) for _ in range(8)]  # churn the allocator
        y = torch.sigmoid(fb.mean(1) @ w)
        loss = torch.nn.functional.binary_cross_entropy(y.clamp(1e-7, 1 - 1e-7), lab)
        loss.backward()
        del tmp
        if k % 10 == 9:
            torch.cuda.synchronize()
            ok = all(bool((cn == float(i)).all()) for i, cn in enumerate(canaries))
            print(k, "canaries ok", ok, "loss", float(loss), "grad finite", bool(torch.isfinite(w.grad).all()), flush=True)
    it.close(); fl.close()
