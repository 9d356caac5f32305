#!/usr/bin/env python
"""Stage times of the cascade's forward pass (64 clips x T=35, bf16): python tools/cascade_times.py"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
sys.argv = sys.argv[:1]
args = bench.parse()
H = bench.Harness(torch.device("cuda:0"), 0, 0, 1)
B, T = 64, 35
cargs = argparse.Namespace(**vars(args))
cargs.model, cargs.batch, cargs.T, cargs.grid, cargs.mode = "gaze_grcn_cascade", B, T, 7, "inference"
model = bench.build_model(cargs, {"rank": 0, "world_size": 1, "allreduce": None}, B, 1, False, "bf16", H.dev)
feat_h, gt_h, _, _ = bench.make_inputs(0, B, T, 7, 49, torch.bfloat16)
model.c3d_input.copy_(feat_h)
model.gt_gazemap.copy_(gt_h)
model.run(train=False, sync=True)
ms, _, _ = H.timed(lambda: model.run(train=False, sync=False), 8, 3)
print(json.dumps({"ms_per_step": ms, "frames_per_s": B * T / (ms * 1e-3), "stage_ms": model.engine.time_stages(model.c3d_input, model.gt_gazemap, iters=3)}))
