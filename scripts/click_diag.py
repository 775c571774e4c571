import sys, os, torch, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/scripts')
import parity_report as PR
from psla_b200.frontend import GpuFbankFrontend
conf = {"num_mel_bins": 128, "target_length": 1056, "mean": PR.MEAN, "std": PR.STD}
dev = torch.device("cuda", 0)
fd, fx = GpuFbankFrontend(conf, device=dev), GpuFbankFrontend(conf, device=dev, exact_clip_mean=True)
for kind in ("click_lowpass_int16", "click_tonal"):
    w = [PR.make(kind, 1000), PR.make(kind, 1001)]
    for lam in (0.37, 0.5):
        a = fd(w, [0], [1], [lam]).cpu()[0, :998] * PR.STD + PR.MEAN
        b = fx(w, [0], [1], [lam]).cpu()[0, :998] * PR.STD + PR.MEAN
        e = (a - b).abs()
        bad = e > 1e-3
        print(kind, "lam", lam, "bad cells", int(bad.sum()), "max", float(e.max()))
        if bad.any():
            print(" per-bin bad counts (nonzero):", {int(i): int(c) for i, c in enumerate(bad.sum(0)) if c})
            fr = bad.sum(1)
            print(" frames with bad cells:", int((fr > 0).sum()), "first few", [int(i) for i in torch.nonzero(fr)[:10, 0]])
            print(" log-mel of bad cells: min %.2f max %.2f median %.2f" % (float(b[bad].min()), float(b[bad].max()), float(b[bad].median())))
            print(" signed mean error on bad cells", float((a - b)[bad].mean()))
