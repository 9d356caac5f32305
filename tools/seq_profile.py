"""Phase timeline of the persistent ConvGRU kernel (CTA 0), from its globaltimer stamps."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from recurrent_gaze_prediction_b200.engine import GazeEngine  # noqa: E402

B = int(os.environ.get("B", 256)); T = 35; Ch = 256
train = os.environ.get("TRAIN", "0") == "1"
eng = GazeEngine("gaze_grcn", B=B, T=T, Ch=Ch, dtype=1, device="cuda", train=train)
eng.params.normal_(0, 0.02); eng.refresh_operands()
eng.Xg.normal_(0, 1.0)
for _ in range(3):
    eng.stage_recurrence()
torch.cuda.synchronize()
buf = (C.c_ulonglong * (32 * 12))()
eng.lib.rgp_convgru_seq_profile(C.addressof(buf), 32 * 12)
ts = np.array(buf[:], dtype=np.int64).reshape(32, 12).astype(np.float64)
names = ["r_gate", "r_issued", "z_issued", "c_gate", "c_issued", "r_full", "r_epi_done", "z_full", "z_epi_done", "c_full", "c_epi_done"]
print("B=%d nsplit=%d train=%d; us relative to the step's r_gate; last column = step period" % (B, eng.nsplit, train))
for t in range(3, 12):
    rel = (ts[t, :11] - ts[t, 0]) / 1e3
    print("t=%2d " % t + " ".join("%s=%.1f" % (n, v) for n, v in zip(names, rel)) + "  period=%.1f" % ((ts[t + 1, 0] - ts[t, 0]) / 1e3))
