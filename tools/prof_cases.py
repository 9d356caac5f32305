#!/usr/bin/env python
"""Short single-purpose runs for `ncu` (one case per invocation, a handful of launches each):
   fwd256   gaze_grcn forward, 256 clips x T=35, bf16 (the bench workload): pack, projection, i2h (im2col), gru_seq, head, loss
   cfg4b8   gaze_lstm 14x14 T=128, 8 clips: i2h + persistent ConvLSTM (lstm_seq_kernel)
   cascade  cascade forward, 32 clips x T=35: up-sampling GEMM, 25-tap top GEMM, topgru_seq, maxout FCs
   train128 gaze_grcn training step, 128 clips
Usage: python tools/prof_cases.py <case> [iters]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.engine import GazeEngine
from recurrent_gaze_prediction_b200.models.model_util import init_parameters, init_parameters_cascade

case = sys.argv[1]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
rng = np.random.RandomState(0)


def tiled(x, B):
    reps = (B + len(x) - 1) // len(x)
    return np.tile(x, (reps,) + (1,) * (x.ndim - 1))[:B]


if case in ("fwd256", "train128"):
    B, T, train = (256, 35, False) if case == "fwd256" else (128, 35, True)
    eng = GazeEngine("gaze_grcn", B=B, T=T, Ch=256, dtype=1, device="cuda", train=train)
    eng.load_tf_params(init_parameters("gaze_grcn", hidden=256, T=T, seed=0))
    c3d = torch.from_numpy(tiled(SY.make_c3d(rng, 8, T), B)).to(torch.bfloat16).cuda()
    gt = torch.from_numpy(tiled(SY.make_gazemaps(rng, 8, T, 49, 49, normalize=True)[0], B)).cuda()
    for _ in range(iters):
        eng.forward(c3d=c3d, gt=gt)
        if train:
            eng.backward()
            eng.optimizer_step()
elif case == "cfg4b8":
    B, T, G = 8, 128, 14
    p = dict(init_parameters("gaze_lstm", hidden=128, T=T, grid=G, seed=0))
    p["out_W"] = np.random.RandomState(7).uniform(-0.1, 0.1, (128, 1)).astype(np.float32)
    eng = GazeEngine("gaze_lstm", B=B, T=T, G=G, Ch=128, dtype=1, device="cuda", head="fc")
    eng.load_tf_params(p)
    emb = torch.from_numpy(SY.make_embedded(rng, B, T, G, 512)).cuda()
    for _ in range(iters):
        eng.forward(emb=emb)
elif case == "refshape":      # the reference's own training shape: small-batch persistent ConvGRU forward + BPTT kernels
    B, T = 28, 42
    eng = GazeEngine("gaze_grcn", B=B, T=T, Ch=128, dtype=1, device="cuda", train=True)
    assert eng.use_seq_small
    eng.load_tf_params(init_parameters("gaze_grcn", hidden=128, T=T, seed=0))
    c3d = torch.from_numpy(tiled(SY.make_c3d(rng, 4, T), B)).to(torch.bfloat16).cuda()
    gt = torch.from_numpy(tiled(SY.make_gazemaps(rng, 4, T, 49, 49, normalize=True)[0], B)).cuda()
    for _ in range(iters):
        eng.forward(c3d=c3d, gt=gt)
        eng.backward()
elif case == "cfg4b64":
    B, T, G = 64, 32, 14
    p = dict(init_parameters("gaze_lstm", hidden=128, T=T, grid=G, seed=0))
    p["out_W"] = np.random.RandomState(7).uniform(-0.1, 0.1, (128, 1)).astype(np.float32)
    eng = GazeEngine("gaze_lstm", B=B, T=T, G=G, Ch=128, dtype=1, device="cuda", head="fc")
    eng.load_tf_params(p)
    emb = torch.from_numpy(SY.make_embedded(rng, B, T, G, 512)).cuda()
    for _ in range(iters):
        eng.forward(emb=emb)
elif case in ("cascade", "cascade2"):
    from recurrent_gaze_prediction_b200.cascade_engine import CascadeEngine
    B, T = (32, 35) if case == "cascade" else (16, 8)
    eng = CascadeEngine(B=B, T=T, dtype=1, device="cuda")
    eng.load_tf_params(init_parameters_cascade(seed=0))
    c3d = torch.from_numpy(tiled(SY.make_c3d(rng, 8, T), B)).to(torch.bfloat16).cuda()
    gt = torch.from_numpy(tiled(SY.make_gazemaps(rng, 8, T, 49, 49)[0], B)).cuda()
    for _ in range(iters):
        eng.forward(c3d=c3d, gt=gt)
else:
    raise SystemExit("unknown case " + case)
torch.cuda.synchronize()
print("case", case, "ok")
