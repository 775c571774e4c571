import sys, os, time, numpy as np, torch
sys.path.insert(0, os.getcwd())
from bench import B_CLIPS, N_MEL, N_SAMPLES, N_WAV, NORM_MEAN, NORM_STD, T_FRAMES
from psla_b200.frontend import GpuFbankFrontend, PackedWaves
dev = torch.device("cuda", 0)
fe = GpuFbankFrontend({"num_mel_bins": N_MEL, "target_length": T_FRAMES, "mean": NORM_MEAN, "std": NORM_STD}, device=dev)
offsets = np.arange(N_WAV, dtype=np.int64) * N_SAMPLES; lengths = np.full(N_WAV, N_SAMPLES, dtype=np.int64)
pk = PackedWaves(0.1 * torch.randn(N_WAV * N_SAMPLES, device=dev), offsets, lengths)
out = torch.empty(B_CLIPS, T_FRAMES, N_MEL, device=dev)
b = np.arange(B_CLIPS); src1 = b.astype(np.int32); src2 = np.where(b % 2 == 0, B_CLIPS + b // 2, -1).astype(np.int32)
lam = np.where(b % 2 == 0, 0.4, 0.0); masks = np.zeros((B_CLIPS, 4), np.int32)
for _ in range(50): fe.forward_packed(pk, src1, src2, lam, masks, out=out)
torch.cuda.synchronize()
ts = []
for rep in range(5):
    t0 = time.perf_counter()
    for _ in range(200): fe.forward_packed(pk, src1, src2, lam, masks, out=out)
    ts.append((time.perf_counter() - t0) / 200 * 1e6); torch.cuda.synchronize()
print(os.environ.get("PSLA_FRONTEND_LIB", "")[-16:], "host us/call", [round(t, 1) for t in ts])
