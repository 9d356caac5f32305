#!/usr/bin/env python
"""Stage breakdown (CUDA events) of one training step: python tools/stage_times.py <model> <B> <T> <Ch> [G]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.engine import GazeEngine, StepGraph
from recurrent_gaze_prediction_b200.models.model_util import init_parameters
model, B, T, Ch = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
rng = np.random.RandomState(0)
eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=1, device="cuda", train=True)
eng.load_tf_params(init_parameters(model, hidden=Ch, T=T, seed=0))
c3d = torch.from_numpy(SY.make_c3d(rng, B, T)).to(torch.bfloat16).cuda()
GH = eng.GH
gt = torch.from_numpy(SY.make_gazemaps(rng, B, T, GH, GH, normalize=True)[0]).cuda()
for _ in range(2):
    eng.forward(c3d=c3d, gt=gt); eng.backward(); eng.optimizer_step()
fw = eng.time_stages(c3d, gt, iters=5)
eng.forward(c3d=c3d, gt=gt)
bw = eng.time_backward(iters=5)
sg = StepGraph(eng, lambda: eng.forward(c3d=c3d, gt=gt), train=True, opt={})
for _ in range(3): sg.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): sg.replay()
e1.record(); torch.cuda.synchronize()
print(json.dumps({"model": model, "B": B, "T": T, "Ch": Ch, "seq_small": bool(getattr(eng, "use_seq_small", False)),
                  "seq_lstm": bool(getattr(eng, "use_seq_lstm", False)), "fwd": fw, "bwd": bw,
                  "graph_step_ms": e0.elapsed_time(e1) / 20, "launches": sg.kernels}))
