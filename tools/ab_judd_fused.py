"""A/B of the two fused AUC-Judd kernels behind rgp_auc_judd_resized (RGP_JUDD_FUSED_KERNEL=old selects the first one):
time per 1024 frames at 49x49 -> 360x480; thresholds / counts / AUCs saved for a bit-level comparison by the caller.
usage: python tools/ab_judd_fused.py out.npz"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recurrent_gaze_prediction_b200 import evaluation_metrics as EM
from recurrent_gaze_prediction_b200 import synthetic as SY
from tools.bench_metrics import ev_time


def main():
    N, H, W = int(os.environ.get("N", 1024)), 360, 480
    rng = np.random.RandomState(0)
    base, _ = SY.make_gazemaps(rng, 8, 8, 49, 49)
    base = base.reshape(64, 49, 49)
    gt = np.tile(base, (N // 64, 1, 1))
    pred = (gt + 0.2 * rng.uniform(size=gt.shape)).astype(np.float32)
    fix = np.tile(SY.make_fixationmaps(rng, base[:, None], H, W).reshape(64, H, W), (N // 64, 1, 1))
    eng = EM.MetricsEngine(chunk=N)
    dev = eng.device
    p64 = eng.to_f64(torch.from_numpy(pred).to(dev), True)
    fx = torch.from_numpy(fix).to(dev)
    torch.manual_seed(0)
    jit = torch.rand(N, H, W, dtype=torch.float64, device=dev)
    maxfix = int(fix.reshape(N, -1).sum(1).max())
    auc, thr, cnt = eng.auc_judd_resized(p64, fx, (H, W), jit, maxfix)
    auc0, thr0, cnt0 = eng.auc_judd_resized(p64, fx, (H, W), None, maxfix)
    torch.cuda.synchronize()
    ms = ev_time(lambda: eng.auc_judd_resized(p64, fx, (H, W), jit, maxfix), iters=10)
    ms0 = ev_time(lambda: eng.auc_judd_resized(p64, fx, (H, W), None, maxfix), iters=10)
    np.savez(sys.argv[1], auc=auc.cpu().numpy(), thr=thr.cpu().numpy(), cnt=cnt.cpu().numpy(), auc0=auc0.cpu().numpy(),
             thr0=thr0.cpu().numpy(), cnt0=cnt0.cpu().numpy())
    print(json.dumps({"kernel": os.environ.get("RGP_JUDD_FUSED_KERNEL", "lean"), "ms_per_%d" % N: ms, "ms_no_jitter": ms0}))


if __name__ == "__main__":
    main()
