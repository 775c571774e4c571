import os, sys, tempfile, torch, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/scripts')
import loader_bench as LB
from psla_b200.dataloaders import AudiosetDataset, FrontendLoader, pcm16_lazy_loader
from psla_b200.frontend import GpuFbankFrontend, PackedWaves
dev = torch.device('cuda', 0)
conf = {"num_mel_bins": 128, "target_length": 1056, "freqm": 48, "timem": 192, "mixup": 0.5, "dataset": "audioset", "mode": "train", "mean": -4.6476, "std": 4.5699, "noise": False}
fe = GpuFbankFrontend(conf, device=dev)
# same prior usage as c5: forward_host first
n_wav=150; N=160000
offs, lens = np.arange(n_wav, dtype=np.int64) * N, np.full(n_wav, N, dtype=np.int64)
pk = PackedWaves(torch.zeros(n_wav*N, dtype=torch.int16).pin_memory(), offs, lens)
b = np.arange(100); src1, src2 = b.astype(np.int32), np.where(b % 2 == 0, 100 + b // 2, -1).astype(np.int32)
f = fe.forward_host(pk, src1, src2, np.full(100, 0.5), None)
torch.cuda.synchronize()
with tempfile.TemporaryDirectory(dir='/dev/shm') as d:
    j, c = LB.make_files(d, 200, seed=0)
    ds = AudiosetDataset(j, conf, label_csv=c, loader=pcm16_lazy_loader, verbose=False)
    sampler = torch.utils.data.WeightedRandomSampler(np.ones(len(ds)), 10 ** 7, replacement=True)
    fl = FrontendLoader.create(ds, fe, batch_size=100, num_workers=15, sampler=sampler, prefetch_factor=2)
    it = iter(fl)
    for k in range(40):
        fb, lab = next(it)
        torch.cuda.synchronize()
        bad = (~torch.isfinite(fb)).sum().item()
        print(k, 'nonfinite', bad, 'fb min/max %.3f %.3f' % (fb.min().item(), fb.max().item()), 'lab min/max %.3f %.3f' % (lab.min().item(), lab.max().item()), flush=True)
        if bad: 
            rows = torch.nonzero(~torch.isfinite(fb).all(dim=2))
            print(' bad rows (clip, frame) first', rows[:5].tolist(), 'count', len(rows))
            break
    it.close(); fl.close()
