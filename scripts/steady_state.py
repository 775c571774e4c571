import sys, os, numpy as np, torch
sys.path.insert(0, '/root/repo')
from psla_b200.frontend import GpuFbankFrontend, PackedWaves
dev = torch.device('cuda', 0)
fe = GpuFbankFrontend({"num_mel_bins": 128, "target_length": 1056, "mean": -4.6476, "std": 4.5699}, device=dev)
hour = 0.1 * torch.randn(57_600_000, device=dev)
n = 160000
for B in (1024, 4096):
    idx = np.arange(B, dtype=np.int64) % 360
    pk = PackedWaves(hour, idx * n, np.full(B, n, dtype=np.int64))
    out = torch.empty(B, 1056, 128, device=dev)
    src1 = np.arange(B, dtype=np.int32)
    for _ in range(2): fe.forward_packed(pk, src1, None, None, None, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fe.forward_packed(pk, src1, None, None, None, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print(B, ms, B / ms * 1e3, 'clips/s')
