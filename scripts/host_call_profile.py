import sys, os, time, cProfile, pstats, numpy as np, torch
sys.path.insert(0, os.getcwd())
from bench import B_CLIPS, N_MEL, N_SAMPLES, N_WAV, NORM_MEAN, NORM_STD, T_FRAMES
from psla_b200 import params as P
from psla_b200.frontend import GpuFbankFrontend, PackedWaves
dev = torch.device("cuda", 0)
fe = GpuFbankFrontend({"num_mel_bins": N_MEL, "target_length": T_FRAMES, "mean": NORM_MEAN, "std": NORM_STD}, device=dev)
offsets = np.arange(N_WAV, dtype=np.int64) * N_SAMPLES; lengths = np.full(N_WAV, N_SAMPLES, dtype=np.int64)
pk = PackedWaves(0.1 * torch.randn(N_WAV * N_SAMPLES, device=dev), offsets, lengths)
out = torch.empty(B_CLIPS, T_FRAMES, N_MEL, device=dev)
b = np.arange(B_CLIPS); src1 = b.astype(np.int32); src2 = np.where(b % 2 == 0, B_CLIPS + b // 2, -1).astype(np.int32)
lam = np.where(b % 2 == 0, 0.4, 0.0); masks = np.zeros((B_CLIPS, 4), np.int32)
for _ in range(20): fe.forward_packed(pk, src1, src2, lam, masks, out=out)
torch.cuda.synchronize()
# time the raw C call
import ctypes
orig = fe.lib.psla_frontend_forward
acc = [0.0, 0]
def timed(*a):
    t0 = time.perf_counter(); r = orig(*a); acc[0] += time.perf_counter() - t0; acc[1] += 1; return r
class L: pass
fe2lib = fe.lib
fe.lib = type("Proxy", (), {"__getattr__": lambda s, n: timed if n == "psla_frontend_forward" else getattr(fe2lib, n)})()
t0 = time.perf_counter()
for _ in range(200): fe.forward_packed(pk, src1, src2, lam, masks, out=out)
t1 = time.perf_counter(); torch.cuda.synchronize()
print("forward_packed us/call %.1f ; C call (incl. ctypes marshalling) us/call %.1f" % ((t1 - t0) / 200 * 1e6, acc[0] / acc[1] * 1e6))
fe.lib = fe2lib
pr = cProfile.Profile(); pr.enable()
for _ in range(200): fe.forward_packed(pk, src1, src2, lam, masks, out=out)
pr.disable(); torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(10)
