"""Ad-hoc: cells on the log floor, fast-log path vs the reference constant."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import psla_oracle as O
from psla_b200.frontend import GpuFbankFrontend
from psla_b200.synth import synth_wave
MEAN, STD = -4.6476, 4.5699
CONF = {"num_mel_bins": 128, "target_length": 1056, "mean": MEAN, "std": STD}
x = synth_wave(1, 160000)
floor = ((torch.log(torch.tensor(O.EPS32, dtype=torch.float32)) - MEAN) / STD)
print("floor const %.9g" % floor.item(), floor.view(torch.int32).item())
for fl in (True, False):
    fe = GpuFbankFrontend(CONF, fast_log=fl)
    o = fe([x]).cpu()
    col = o[0, :998, 3]
    print("fast_log", fl, "unique values in mel row 3:", [(float(v), int(torch.tensor(v).view(torch.int32))) for v in col.unique()[:5]])
    z = fe([torch.zeros(160000)]).cpu()
    print("   zeros clip unique:", [(float(v), int(torch.tensor(v).view(torch.int32))) for v in z[0, :998].unique()[:5]])
