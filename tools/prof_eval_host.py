"""cProfile of MetricsEngine.evaluate_batch on 1024 cascade-sized maps at 360x480 (host-side cost of the evaluation call)."""
import cProfile
import os
import pstats
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from recurrent_gaze_prediction_b200 import evaluation_metrics as EM
from recurrent_gaze_prediction_b200 import synthetic as SY


def main():
    N, H, W = 1024, 360, 480
    rng = np.random.RandomState(0)
    base, _ = SY.make_gazemaps(rng, 8, 8, 49, 49)
    base = base.reshape(64, 49, 49)
    gt = np.tile(base, (N // 64, 1, 1))
    pred = (gt + 0.2 * rng.uniform(size=gt.shape)).astype(np.float32)
    fix = np.tile(SY.make_fixationmaps(rng, base[:, None], H, W).reshape(64, H, W), (N // 64, 1, 1))
    union = (fix[:10] > 0).sum(0).astype(np.float64)
    eng = EM.MetricsEngine(chunk=256)
    mets = ('sim', 'cc', 'NSS', 'AUC_Judd', 'AUC_Borji', 'AUC_shuffled')
    for mode in ("device", "numpy"):
        kw = dict(rng="device") if mode == "device" else {}
        eng.evaluate_batch(pred[:64], gt[:64], fix[:64], union=union, metrics=mets, **kw)
        torch.cuda.synchronize()
        pr = cProfile.Profile()
        pr.enable()
        eng.evaluate_batch(pred, gt, fix, union=union, metrics=mets, **kw)
        torch.cuda.synchronize()
        pr.disable()
        print("=== rng =", mode)
        pstats.Stats(pr).sort_stats("cumulative").print_stats(22)


if __name__ == "__main__":
    main()
