#!/usr/bin/env python
"""Phase timeline of the persistent ConvLSTM kernel (rgp_convlstm_seq_profile): CTA 0, steps 4..31, microseconds from the
step's dependency-satisfied stamp.  Usage: python tools/lstm_seq_timeline.py [B] [G] [T]"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from recurrent_gaze_prediction_b200.engine import GazeEngine
from recurrent_gaze_prediction_b200.models.model_util import init_parameters
from recurrent_gaze_prediction_b200 import synthetic as SY

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
G = int(sys.argv[2]) if len(sys.argv) > 2 else 14
T = int(sys.argv[3]) if len(sys.argv) > 3 else 128
p = dict(init_parameters("gaze_lstm", hidden=128, T=T, grid=G, seed=0))
p["out_W"] = np.random.RandomState(7).uniform(-0.1, 0.1, (128, 1)).astype(np.float32)
eng = GazeEngine("gaze_lstm", B=B, T=T, G=G, Ch=128, dtype=1, device="cuda", head="fc")
assert eng.use_seq_lstm
eng.load_tf_params(p)
emb = torch.from_numpy(SY.make_embedded(np.random.RandomState(0), B, T, G, 512)).cuda()
for _ in range(3):
    eng.forward(emb=emb)
torch.cuda.synchronize()
st = eng.time_stages(emb, torch.zeros(B, T, G, G, device="cuda"), iters=3, emb=True)
buf = (C.c_ulonglong * (32 * 8))()
eng.lib.rgp_convlstm_seq_profile(C.addressof(buf), 32 * 8)
ts = np.array(list(buf), dtype=np.float64).reshape(32, 8)
names = ["deps ok", "slab TMA issued", "slab landed", "MMAs issued", "acc seen", "gates+store", "published"]
rel = (ts[4:31, :7] - ts[4:31, :1]) / 1e3
period = np.diff(ts[4:32, 0]) / 1e3
out = {"B": B, "G": G, "T": T, "recurrence_ms": st["recurrence"], "us_per_step": st["recurrence"] * 1e3 / T,
       "period_us_median": float(np.median(period)),
       "timeline_us_from_deps_ok": {n: float(np.median(rel[:, i])) for i, n in enumerate(names)}, "stage_ms": st}
print(json.dumps(out))
