"""First-contact diagnostics on a B200 (run under gpurun): prints error statistics of the CUDA frontend
against the live oracle instead of just asserting, and a quick CUDA-event timing."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import psla_oracle as O  # noqa: E402  (checker only)
from psla_b200.frontend import GpuFbankFrontend, pack_waveforms  # noqa: E402
from psla_b200.synth import synth_wave  # noqa: E402

torch.set_num_threads(8)
CONF = {"num_mel_bins": 128, "target_length": 1056, "mean": -4.6476, "std": 4.5699}
res = {}
print(torch.cuda.get_device_name(0), torch.version.cuda)

fe = GpuFbankFrontend(CONF)
x = synth_wave(0, 160000)
for fast in (False, True):
    fe.set_fast_log(fast)
    out = fe([x]).cpu()
    ref = O.batch_torch([x], [0], [-1], [0.0], [(0, 0, 0, 0)], 1056, -4.6476, 4.5699)
    ref64 = O.batch_torch([x], [0], [-1], [0.0], [(0, 0, 0, 0)], 1056, -4.6476, 4.5699, dtype=torch.float64)
    d = (out - ref).abs()
    print(f"C1 fast_log={fast}: max|gpu-ref32|={d.max().item():.3e} (x std: {d.max().item()*4.5699:.3e} in log-mel units) "
          f"max|gpu-ref64|={(out.double()-ref64).abs().max().item():.3e} max|ref32-ref64|={(ref.double()-ref64).abs().max().item():.3e}")
    fill = ((torch.zeros(1) - (-4.6476)) / 4.5699).item()
    print("  pad rows exact:", bool((out[0, 998:] == fill).all()), "nan:", bool(torch.isnan(out).any()))
    res[f"c1_fast{int(fast)}"] = d.max().item() * 4.5699

# batch with mixup / masks
waves = [synth_wave(s, n, k) for s, n, k in [(1, 160000, "gauss"), (2, 160000, "am"), (3, 100000, "int16"), (4, 200000, "gauss"), (5, 5000, "gauss")]]
src1 = [0, 1, 2, 0, 3, 4, 2]
src2 = [-1, 0, 3, 2, 1, -1, 2]
lam = [0.0, 0.3, 0.55, 0.7, 0.45, 0.0, 0.5]
masks = [(0, 0, 0, 0), (10, 20, 100, 150), (0, 47, 900, 156), (127, 1, 0, 1), (50, 0, 500, 0), (3, 5, 2, 3), (100, 28, 1055, 1)]
out = fe(waves, src1, src2, lam, masks).cpu()
ref = O.batch_torch(waves, src1, src2, lam, masks, 1056, -4.6476, 4.5699)
d = (out - ref).abs() * 4.5699
print("batch: per-clip max err (log-mel units):", [f"{v:.2e}" for v in d.amax((1, 2)).tolist()])
print("batch: fill cells identical:", bool(((out == fill) == (ref == fill)).all()))
res["batch"] = d.max().item()

# timing: 100-clip AudioSet-shaped batch, device resident
B = 100
g = torch.Generator().manual_seed(0)
big = [0.1 * torch.randn(1, 160000, generator=g) for _ in range(150)]
packed = pack_waveforms(big, device="cuda")
s1 = np.arange(B, dtype=np.int32)
s2 = np.where(np.arange(B) % 2 == 0, 100 + np.arange(B) // 2, -1).astype(np.int32)
lm = np.full(B, 0.5)
mk = np.tile(np.array([[20, 24, 300, 96]], np.int32), (B, 1))
outb = torch.empty(B, 1056, 128, device="cuda")
for fast in (False, True):
    fe.set_fast_log(fast)
    for _ in range(5):
        fe.forward_packed(packed, s1, s2, lm, mk, out=outb)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 50
    e0.record()
    for _ in range(n):
        fe.forward_packed(packed, s1, s2, lm, mk, out=outb)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print(f"train batch B=100 fast_log={fast}: {ms:.4f} ms/step -> {B/ms*1e3:.3e} clips/s")
    res[f"train100_fast{int(fast)}_clips_s"] = B / ms * 1e3
# eval, bigger batch
for Bn in (100, 1000):
    idx = (np.arange(Bn) % 150).astype(np.int32)
    outn = torch.empty(Bn, 1056, 128, device="cuda")
    for _ in range(3):
        fe.forward_packed(packed, idx, out=outn)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        fe.forward_packed(packed, idx, out=outn)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"eval B={Bn}: {ms:.4f} ms/step -> {Bn/ms*1e3:.3e} clips/s")
    res[f"eval{Bn}_clips_s"] = Bn / ms * 1e3
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w"), indent=1)

# ---- C2 error anatomy (100 clips): where do cells beyond 1e-3 come from? ----
fe.set_fast_log(True)
wv = [synth_wave(s, 160000, ("gauss", "am", "int16")[s % 3]) for s in range(150)]
s1 = list(range(100)); s2 = [100 + b // 2 if b % 2 == 0 else -1 for b in range(100)]
ld = np.random.default_rng(0).beta(10, 10, 50); lm2 = [float(ld[b // 2]) if b % 2 == 0 else 0.0 for b in range(100)]
torch.manual_seed(0); mk2 = [O.draw_mask_offsets(48, 192, 128, 1056) for _ in range(100)]
o = fe(wv, s1, s2, lm2, mk2).cpu()
r32 = O.batch_torch(wv, s1, s2, lm2, mk2, 1056, -4.6476, 4.5699)
r64 = O.batch_torch(wv, s1, s2, lm2, mk2, 1056, -4.6476, 4.5699, dtype=torch.float64)
S = 4.5699
e = (o - r32).abs() * S; eo = (o.double() - r64).abs() * S; er = (r32.double() - r64).abs() * S
print(f"C2: n={e.numel()} max|gpu-ref32|={e.max():.3e} cells>1e-3: {(e > 1e-3).sum().item()}  >5e-4: {(e > 5e-4).sum().item()}")
print(f"C2: max|gpu-ref64|={eo.max():.3e} cells>1e-3: {(eo > 1e-3).sum().item()} ; max|ref32-ref64|={er.max():.3e} cells>1e-3: {(er > 1e-3).sum().item()}")
print(f"C2: rms gpu-ref64={eo.pow(2).mean().sqrt():.3e} rms ref32-ref64={er.pow(2).mean().sqrt():.3e}")
idx = (e > 1e-3).nonzero()
for b_, t_, j_ in idx[:12].tolist():
    print(f"   cell clip {b_} frame {t_} mel {j_}: logmel64={r64[b_, t_, j_] * S - 4.6476:.3f} gpu_err64={eo[b_, t_, j_]:.2e} ref_err64={er[b_, t_, j_]:.2e}")
res["c2_max_vs_ref32"] = e.max().item(); res["c2_gt1e3"] = int((e > 1e-3).sum()); res["c2_max_gpu64"] = eo.max().item(); res["c2_max_ref64"] = er.max().item()
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "gpu_check.json"), "w"), indent=1)
