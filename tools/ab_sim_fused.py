"""A/B of the two fused sim/cc/NSS kernels behind rgp_sim_cc_nss_resized (RGP_SIM_FUSED_KERNEL=old selects the first one):
time per 1024 frames at 49x49 -> 360x480 and the statistics rows, saved for a bit-level comparison by the caller.
usage: python tools/ab_sim_fused.py out.npy [noise]"""
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
    noise = len(sys.argv) > 2 and sys.argv[2] == "noise"
    rng = np.random.RandomState(0)
    base, _ = SY.make_gazemaps(rng, 8, 8, 49, 49)
    base = base.reshape(64, 49, 49)
    gt = np.tile(base, (N // 64, 1, 1))
    pred = (gt + 0.2 * rng.uniform(size=gt.shape)).astype(np.float32)
    if noise:                                  # worst case for the clip-free windows: white noise in both maps
        pred = rng.uniform(size=gt.shape).astype(np.float32)
        gt = rng.uniform(size=gt.shape).astype(np.float32)
    fix = np.tile(SY.make_fixationmaps(rng, base[:, None], H, W).reshape(64, H, W), (N // 64, 1, 1))
    eng = EM.MetricsEngine(chunk=N)
    dev = eng.device
    p64 = eng.to_f64(torch.from_numpy(pred).to(dev), True)
    g64 = eng.to_f64(torch.from_numpy(gt).to(dev), False)
    fx = torch.from_numpy(fix).to(dev)
    out = eng.sim_cc_nss_resized(p64, g64, fx, (H, W))
    torch.cuda.synchronize()
    ms = ev_time(lambda: eng.sim_cc_nss_resized(p64, g64, fx, (H, W)), iters=10)
    np.save(sys.argv[1], out.cpu().numpy())
    print(json.dumps({"kernel": os.environ.get("RGP_SIM_FUSED_KERNEL", "lean"), "noise": noise, "ms_per_%d" % N: ms,
                      "GBps_on_9B_per_px": N * H * W * 9 / ms / 1e6}))


if __name__ == "__main__":
    main()
