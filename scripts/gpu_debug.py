"""Ad-hoc GPU diagnostics (run under gpurun): exact-mean vs fused-mean on the C2 batch."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import psla_oracle as O
from psla_b200.frontend import GpuFbankFrontend
from psla_b200.synth import synth_wave
MEAN, STD = -4.6476, 4.5699
CONF = {"num_mel_bins": 128, "target_length": 1056, "mean": MEAN, "std": STD}
B = 100
waves = [synth_wave(s, 160000, ("gauss", "am", "int16")[s % 3]) for s in range(150)]
src1 = list(range(B)); src2 = [100 + b // 2 if b % 2 == 0 else -1 for b in range(B)]
ld = np.random.default_rng(0).beta(10, 10, 50); lam = [float(ld[b // 2]) if b % 2 == 0 else 0.0 for b in range(B)]
torch.manual_seed(0); masks = [O.draw_mask_offsets(48, 192, 128, 1056) for _ in range(B)]
fill = ((torch.zeros(1) - MEAN) / STD).item()
outs = {}
for exact in (False, True):
    fe = GpuFbankFrontend(CONF, exact_clip_mean=exact)
    for rep in range(3):
        o = fe(waves, src1, src2, lam, masks).cpu()
        torch.cuda.synchronize()
        print("exact", exact, "rep", rep, "nan", int(torch.isnan(o).sum()), "fillcells", int((o == fill).sum()), "launches", fe.launch_count)
    outs[exact] = o
d = (outs[True] - outs[False]).abs()
print("max diff exact-fused", d.max().item())
bad = ((outs[True] == fill) != (outs[False] == fill)).nonzero()
print("fill mismatch cells", len(bad), bad[:10].tolist())
if len(bad):
    b_, t_, j_ = bad[0].tolist()
    print("clip", b_, "mask", masks[b_], "src2", src2[b_], "row", t_, outs[True][b_, t_, :8], outs[False][b_, t_, :8])
    rows = sorted(set(bad[:, 1].tolist()))
    print("clips:", sorted(set(bad[:, 0].tolist()))[:20], "rows:", rows[:40])
