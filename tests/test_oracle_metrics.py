"""The metrics oracle against goldens minted by RUNNING the reference's evaluation_metrics.py
(oracle/gen_golden_metrics.py).  CPU only."""
import os

import numpy as np
import pytest

from oracle import metrics_oracle as MO

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "metrics_ref.npz"))
JITTER_SEED0 = 5000


def _eq(a, b, tol=1e-12):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert np.array_equal(np.isnan(a), np.isnan(b))
    m = ~np.isnan(a)
    assert np.max(np.abs(a[m] - b[m]), initial=0.0) <= tol, np.max(np.abs(a[m] - b[m]))


@pytest.mark.parametrize("tag", ["s49", "vas"])
def test_metrics_oracle_matches_reference_run(tag):
    pred, gt, fix = G[tag + "_pred"].astype(np.float64), G[tag + "_gt"].astype(np.float64), G[tag + "_fix"]
    union = G[tag + "_union"]
    N, H, W = pred.shape
    got = {k: [] for k in ("sim", "cc", "judd", "judd_nojit", "borji", "shuf")}
    for i in range(N):
        p = MO.normalize_range(pred[i])
        n_fix = int((fix[i] > 0.5).sum())
        jit = np.random.RandomState(JITTER_SEED0 + i).rand(H, W)
        ridx = G[tag + "_ridx"][i][:max(n_fix, 1)].astype(np.int64)
        sidx = G[tag + "_sidx"][i]
        sidx = sidx[(sidx >= 0).all(axis=1)].astype(np.int64)
        got["sim"].append(MO.similarity(gt[i], p))
        got["cc"].append(MO.cc(gt[i], p))
        got["judd"].append(MO.AUC_Judd(fix[i], p, jitter_values=jit))
        got["judd_nojit"].append(MO.AUC_Judd(fix[i], p, jitter=False))
        got["borji"].append(MO.AUC_Borji(fix[i], p, rand_index=ridx))
        got["shuf"].append(MO.AUC_shuffled(fix[i], p, union, rand_index=sidx))
    for k in got:
        _eq(got[k], G[tag + "_ref_" + k])
    assert np.isnan(got["judd"][5]) and np.isnan(got["borji"][5])


def test_global_rng_draw_order_matches_reference_contract():
    """With jitter_values/rand_index omitted the oracle draws from numpy's global RNG exactly
    like the reference (random.rand(H,W); random.randint(0,n_pix,[n_fix,n_rep]))."""
    pred, fix = MO.normalize_range(G["s49_pred"][0].astype(np.float64)), G["s49_fix"][0]
    np.random.seed(3)
    a = MO.AUC_Judd(fix, pred)
    np.random.seed(3)
    b = MO.AUC_Judd(fix, pred, jitter_values=np.random.rand(49, 49))
    assert a == b
    np.random.seed(4)
    a = MO.AUC_Borji(fix, pred)
    np.random.seed(4)
    n_fix = int((fix > 0.5).sum())
    b = MO.AUC_Borji(fix, pred, rand_index=np.random.randint(0, 49 * 49, [n_fix, 100]))
    assert a == b


def test_helpers_match_reference():
    _eq(MO.normalize_range(G["nr_in"]), G["nr_ref"], 0.0)
    assert np.array_equal(MO.resize_onehot_tensor_sparse(G["onehot_in"], (49, 49)).astype(np.uint8), G["onehot_ref"])


def test_resize_bicubic_properties():
    rng = np.random.RandomState(0)
    x = rng.uniform(size=(49, 49))
    assert np.array_equal(MO.resize_bicubic(x, (49, 49)), x)
    y = MO.resize_bicubic(x, (360, 480), mode="nearest")
    assert y.shape == (360, 480) and y.min() >= x.min() and y.max() <= x.max()
    c = np.full((7, 7), 0.25)
    assert np.allclose(MO.resize_bicubic(c, (21, 28), mode="nearest"), 0.25, atol=1e-15)
    # linear ramps are reproduced by Catmull-Rom away from the border
    r = np.tile(np.arange(20.0), (20, 1))
    z = MO.resize_bicubic(r, (40, 40), mode="nearest")
    cs = (np.arange(40) + 0.5) * 0.5 - 0.5
    assert np.allclose(z[10, 4:-4], cs[4:-4], atol=1e-12)


def test_nss_definition():
    rng = np.random.RandomState(1)
    s = rng.uniform(size=(30, 40))
    f = np.zeros((30, 40)); f[3, 4] = f[10, 11] = 1
    z = (s - s.mean()) / s.std()
    assert abs(MO.nss(f, s) - (z[3, 4] + z[10, 11]) / 2) < 1e-14
    assert np.isnan(MO.nss(np.zeros((30, 40)), s))


def test_saliency_score_nan_propagates_and_needs_ten_frames():
    pred, gt, fix = G["s49_pred"].astype(np.float64), G["s49_gt"].astype(np.float64), G["s49_fix"]
    np.random.seed(0)
    assert np.isnan(MO.saliency_score("AUC_Borji", list(pred), list(gt), list(fix)))  # frame 5 has no fixation
    with pytest.raises(AssertionError):
        MO.saliency_score("sim", list(pred[:5]), list(gt[:5]), list(fix[:5]))
    v = MO.saliency_score("sim", list(pred), list(gt), list(fix))
    assert 0 < v < 1
