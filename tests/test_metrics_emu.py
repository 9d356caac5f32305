"""Host side of the metric path (recurrent_gaze_prediction_b200/evaluation_metrics.py) over the NumPy emulation
of the C-ABI, against (1) goldens minted by running the reference's evaluation_metrics.py and (2) the oracle.
CPU only; the CUDA kernels themselves are checked in test_gpu_metrics.py."""
import os

import numpy as np
import pytest

from oracle import metrics_oracle as MO
from recurrent_gaze_prediction_b200 import evaluation_metrics as EM
from emu_lib import EmuLib

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "metrics_ref.npz"))
JITTER_SEED0 = 5000


def golden_case(tag):
    pred, gt, fix = G[tag + "_pred"], G[tag + "_gt"], G[tag + "_fix"]
    N, H, W = pred.shape
    nfix = (fix > 0.5).reshape(N, -1).sum(1)
    jit = np.stack([np.random.RandomState(JITTER_SEED0 + i).rand(H, W) for i in range(N)])
    ridx = [G[tag + "_ridx"][i][:nfix[i]].astype(np.int64) for i in range(N)]
    sidx = []
    for i in range(N):
        s = G[tag + "_sidx"][i]
        sidx.append(s[(s >= 0).all(axis=1)].astype(np.int64) if nfix[i] else np.zeros((0, 100), np.int64))
    return pred, gt, fix, G[tag + "_union"], jit, ridx, sidx


def check_against_golden(eng, tag, tol=1e-12):
    pred, gt, fix, union, jit, ridx, sidx = golden_case(tag)
    r = eng.evaluate_batch(pred, gt, fix, union=union, jitter_values=jit, rand_index=ridx, shuf_index=sidx)
    r0 = eng.evaluate_batch(pred, gt, fix, metrics=("AUC_Judd",), jitter=False)
    for ours, ref in (("sim", "sim"), ("cc", "cc"), ("AUC_Judd", "judd"), ("AUC_Borji", "borji"),
                      ("AUC_shuffled", "shuf")):
        a, b = r[ours], G[tag + "_ref_" + ref]
        assert np.array_equal(np.isnan(a), np.isnan(b)), ours
        m = ~np.isnan(b)
        assert np.abs(a[m] - b[m]).max() <= tol, (ours, np.abs(a[m] - b[m]).max())
    b = G[tag + "_ref_judd_nojit"]
    m = ~np.isnan(b)
    assert np.abs(r0["AUC_Judd"][m] - b[m]).max() <= tol
    # NSS has no reference: the oracle's definition
    for i in range(len(pred)):
        want = MO.nss(fix[i], MO.normalize_range(pred[i].astype(np.float64)))
        assert (np.isnan(want) and np.isnan(r["NSS"][i])) or abs(want - r["NSS"][i]) <= 1e-9


@pytest.mark.parametrize("tag", ["s49", "vas"])
def test_batch_path_matches_reference_goldens(tag):
    check_against_golden(EM.MetricsEngine(device="cpu", lib=EmuLib()), tag)


def test_reference_style_functions_and_rng_order():
    eng = EM.MetricsEngine(device="cpu", lib=EmuLib())
    pred, gt, fix = G["s49_pred"], G["s49_gt"], G["s49_fix"]
    p0 = MO.normalize_range(pred[0].astype(np.float64))
    # global-RNG draws happen where the reference draws them
    np.random.seed(3)
    a = EM.AUC_Judd(fix[0], p0, engine=eng)
    np.random.seed(3)
    assert a == MO.AUC_Judd(fix[0], p0)
    np.random.seed(4)
    a = EM.AUC_Borji(fix[0], p0, engine=eng)
    np.random.seed(4)
    assert abs(a - MO.AUC_Borji(fix[0], p0)) <= 1e-12
    np.random.seed(5)
    a = EM.AUC_shuffled(fix[0], p0, G["s49_union"], engine=eng)
    np.random.seed(5)
    assert abs(a - MO.AUC_shuffled(fix[0], p0, G["s49_union"])) <= 1e-12
    assert abs(EM.similarity(gt[0], p0, engine=eng) - MO.similarity(gt[0], p0)) <= 1e-12
    assert abs(EM.cc(gt[0], p0, engine=eng) - MO.cc(gt[0], p0)) <= 1e-12
    # no fixation -> NaN ; unknown metric / shape mismatch -> ValueError (evaluation_metrics.py:71-73,196,270)
    assert np.isnan(EM.AUC_Judd(np.zeros((49, 49)), p0, engine=eng))
    assert np.isnan(EM.saliency_score_single("AUC_Borji", pred[5], gt[5], fix[5], engine=eng))
    with pytest.raises(ValueError):
        EM.saliency_score_single("bogus", pred[0], gt[0], fix[0], engine=eng)
    with pytest.raises(ValueError):
        EM.AUC_shuffled(fix[0], p0, np.zeros((3, 3)), engine=eng)
    with pytest.raises(ValueError):
        EM.saliency_score_single("AUC_shuffled", pred[0], gt[0], fix[0], engine=eng)
    assert EM.AVAILABLE_METRICS == ['sim', 'cc', 'AUC_shuffled', 'AUC_Borji']


def test_saliency_score_single_with_resize_matches_oracle():
    """49x49 model maps against a 60x80 fixation map: prep -> bicubic resize -> metric, as evaluation_metrics.py:239-272."""
    eng = EM.MetricsEngine(device="cpu", lib=EmuLib())
    pred, gt, fix = G["s49_pred"], G["s49_gt"], G["vas_fix"]
    union = G["vas_union"]
    for i in (0, 3):
        n_fix = int((fix[i] > 0.5).sum())
        jit = np.random.RandomState(i).rand(60, 80)
        ridx = np.random.RandomState(50 + i).randint(0, 4800, [n_fix, 100])
        for metric, kw in (("sim", {}), ("cc", {}), ("NSS", {}), ("AUC_Judd", dict(jitter_values=jit)),
                           ("AUC_Borji", dict(rand_index=ridx))):
            a = EM.saliency_score_single(metric, pred[i], gt[i], fix[i], union, engine=eng, **kw)
            b = MO.saliency_score_single(metric, pred[i].astype(np.float64), gt[i], fix[i], union, **kw)
            assert abs(a - b) <= 1e-10, (metric, a, b)


def test_saliency_score_mean_and_nan_propagation():
    eng = EM.MetricsEngine(device="cpu", lib=EmuLib())
    pred, gt, fix = G["s49_pred"], G["s49_gt"], G["s49_fix"]
    np.random.seed(0)
    a = EM.saliency_score("sim", list(pred), list(gt), list(fix), engine=eng)
    np.random.seed(0)
    b = MO.saliency_score("sim", list(pred.astype(np.float64)), list(gt), list(fix))
    assert abs(a - b) <= 1e-12
    np.random.seed(1)
    assert np.isnan(EM.saliency_score("AUC_Borji", list(pred), list(gt), list(fix), engine=eng))   # frame 5: no fixation
    keep = [i for i in range(12) if i != 5]
    np.random.seed(2)
    a = EM.saliency_score("AUC_shuffled", list(pred[keep]), list(gt[keep]), list(fix[keep]), engine=eng)
    np.random.seed(2)
    b = MO.saliency_score("AUC_shuffled", list(pred[keep].astype(np.float64)), list(gt[keep]), list(fix[keep]))
    assert abs(a - b) <= 1e-12
    with pytest.raises(AssertionError):
        EM.saliency_score("sim", list(pred[:5]), list(gt[:5]), list(fix[:5]), engine=eng)


def test_resize_onehot_matches_reference_golden():
    assert np.array_equal(EM.resize_onehot_tensor_sparse(G["onehot_in"], (49, 49)).astype(np.uint8), G["onehot_ref"])
    assert np.array_equal(EM.normalize_range(G["nr_in"]), G["nr_ref"])


def test_fused_and_two_step_host_paths_agree():
    """evaluate_batch routes below-resolution maps to the rgp_*_resized entry points (fused="always") or to
    rgp_resize_bicubic + the plain ones; over the emulation both must give the same numbers."""
    rng = np.random.RandomState(4)
    N, H, W = 3, 30, 44
    pred = rng.uniform(size=(N, 7, 7)).astype(np.float32)
    gt = rng.uniform(size=(N, 7, 7)).astype(np.float32)
    fix = (rng.uniform(size=(N, H, W)) > 0.99).astype(np.uint8)
    fix[1] = 0
    union = (rng.uniform(size=(H, W)) > 0.9).astype(np.float64)
    jit = rng.rand(N, H, W)
    nfix = fix.reshape(N, -1).sum(1)
    ridx = [rng.randint(0, H * W, [int(n), 20]) for n in nfix]
    lib = EmuLib()
    calls = []
    for fused in ("always", False):
        eng = EM.MetricsEngine(device="cpu", lib=lib, fused=fused)
        np.random.seed(1)
        calls.append(eng.evaluate_batch(pred, gt, fix, union=union, jitter_values=jit, rand_index=ridx, n_rep=20))
    for k in calls[0]:
        assert np.array_equal(calls[0][k], calls[1][k], equal_nan=True), k
    assert np.isnan(calls[0]["AUC_Judd"][1]) and not np.isnan(calls[0]["AUC_Judd"][0])
    want = MO.saliency_score_single("AUC_Judd", pred[0].astype(np.float64), gt[0], fix[0], jitter_values=jit[0])
    assert abs(calls[0]["AUC_Judd"][0] - want) <= 1e-12


def test_sparse_fixation_maps_are_accepted():
    """evaluation_metrics.py:240: scipy.sparse fixation maps (what the reference's loaders emit,
    crc_input_data_seq.py:277-321) give the same scores as their dense form."""
    import scipy.sparse as sp
    eng = EM.MetricsEngine(device="cpu", lib=EmuLib())
    pred, gt, fix, union = G["s49_pred"], G["s49_gt"], G["s49_fix"], G["s49_union"]
    i = 0
    ridx = np.random.RandomState(3).randint(0, 2401, [int((fix[i] > 0.5).sum()), 100])
    for metric, kw in (("sim", {}), ("NSS", {}), ("AUC_Borji", dict(rand_index=ridx))):
        a = EM.saliency_score_single(metric, pred[i], gt[i], sp.csr_matrix(fix[i]), sp.csr_matrix(union), engine=eng, **kw)
        b = EM.saliency_score_single(metric, pred[i], gt[i], fix[i], union, engine=eng, **kw)
        assert a == b, metric
    keep = [k for k in range(12) if k != 5]
    np.random.seed(4)
    a = EM.saliency_score("cc", list(pred[keep]), list(gt[keep]), [sp.csr_matrix(fix[k]) for k in keep], engine=eng)
    np.random.seed(4)
    b = EM.saliency_score("cc", list(pred[keep]), list(gt[keep]), list(fix[keep]), engine=eng)
    assert a == b


def test_device_rng_mode_is_statistically_equivalent():
    """rng="device" (opt-in): jitter / Borji negatives / shuffled picks from the counter-based device generator instead of
    NumPy's global stream - the deterministic metrics are untouched, the sampled AUCs agree with the reference-stream path
    to sampling noise, and the call no longer consumes the global RNG."""
    from recurrent_gaze_prediction_b200 import evaluation_metrics as EM
    from recurrent_gaze_prediction_b200 import synthetic as SY
    rng = np.random.RandomState(3)
    gt, _ = SY.make_gazemaps(rng, 6, 2, 49, 49)
    gt = gt.reshape(12, 49, 49)
    pred = (gt + 0.3 * rng.uniform(size=gt.shape)).astype(np.float32)
    fix = SY.make_fixationmaps(rng, gt[:, None], 49, 49).reshape(12, 49, 49)
    union = (fix[:6].sum(axis=0) > 0).astype(np.uint8)
    eng = EM.MetricsEngine(device="cpu", lib=EmuLib())
    np.random.seed(5)
    a = eng.evaluate_batch(pred, gt, fix, union=union)
    state = np.random.get_state()[1].copy()
    b = eng.evaluate_batch(pred, gt, fix, union=union, rng="device", seed=9)
    assert np.array_equal(np.random.get_state()[1], state)             # the global stream was not consumed
    c = eng.evaluate_batch(pred, gt, fix, union=union, rng="device", seed=9)
    for m in ("sim", "cc", "NSS"):
        assert np.array_equal(a[m], b[m], equal_nan=True)
    for m in ("AUC_Judd", "AUC_Borji", "AUC_shuffled"):
        assert np.array_equal(b[m], c[m], equal_nan=True)               # reproducible from the seed
        ok = np.isfinite(a[m])
        assert np.abs(a[m][ok] - b[m][ok]).max() < 0.08, m              # sampling noise of 100 repetitions
    with pytest.raises(ValueError):
        eng.evaluate_batch(pred, gt, fix, union=union, rng="philox")
