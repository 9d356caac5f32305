"""GPU parity of the saliency-metric kernels (csrc/metrics.cu, through the C-ABI) against the reference goldens and
the float64 oracle.  Integer contracts (AUC-Judd rank counts, Borji / shuffled tp/fp counts) and the thresholds are
BIT-EXACT; sim / cc / NSS are reductions whose summation order differs from NumPy's pairwise sum: 1e-10 absolute
(the north star asks for +-0.005)."""
import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO
from recurrent_gaze_prediction_b200 import evaluation_metrics as EM
from recurrent_gaze_prediction_b200 import synthetic as SY
from test_metrics_emu import G, check_against_golden, golden_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    return EM.MetricsEngine(device="cuda")


@pytest.mark.parametrize("tag", ["s49", "vas"])
def test_kernels_match_reference_goldens(eng, tag):
    check_against_golden(eng, tag, tol=1e-12)


@pytest.mark.parametrize("tag", ["s49", "vas"])
def test_integer_contracts_bit_exact(eng, tag):
    pred, gt, fix, union, jit, ridx, sidx = golden_case(tag)
    r = eng.evaluate_batch(pred, gt, fix, union=union, jitter_values=jit, rand_index=ridx, shuf_index=sidx,
                           return_counts=True)
    N = len(pred)
    for i in range(N):
        F = (fix[i] > 0.5).ravel()
        n_fix = int(F.sum())
        if n_fix == 0:
            assert (r["AUC_Judd_counts"][i] == -1).all()
            continue
        p = MO.normalize_range(pred[i].astype(np.float64))
        S = MO.normalize_range(p + jit[i] * 1e-7).ravel()
        thr, above = MO.auc_judd_counts(S, F)
        assert np.array_equal(r["AUC_Judd_thresholds"][i][:n_fix], thr)          # float64 bit pattern
        assert np.array_equal(r["AUC_Judd_counts"][i][:n_fix], above)
        assert (r["AUC_Judd_counts"][i][n_fix:] == -1).all()
        S = MO.normalize_range(p).ravel()
        for name, index in (("AUC_Borji", ridx[i]), ("AUC_shuffled", sidx[i])):
            for rep, (thresholds, tpc, fpc) in enumerate(MO.borji_counts(S[F], S[index], 0.1)):
                L = len(thresholds)
                assert np.array_equal(r[name + "_tp"][i, rep, :L], tpc[::-1]), (name, i, rep)
                assert np.array_equal(r[name + "_fp"][i, rep, :L], fpc[::-1]), (name, i, rep)
                assert (r[name + "_tp"][i, rep, L:] == -1).all()


def test_resize_bicubic_bit_exact_vs_oracle(eng):
    rng = np.random.RandomState(0)
    x = rng.uniform(size=(5, 49, 49))
    for shape, mode in (((360, 480), "constant"), ((360, 480), "nearest"), ((60, 80), "constant"), ((20, 31), "nearest")):
        y = eng.resize(torch.from_numpy(x).cuda(), shape, mode).cpu().numpy()
        for i in range(len(x)):
            assert np.array_equal(y[i], MO.resize_bicubic(x[i], shape, mode)), (shape, mode, i)
    big = rng.uniform(size=(2, 200, 300))          # source too large for the shared-memory path
    y = eng.resize(torch.from_numpy(big).cuda(), (360, 480), "constant").cpu().numpy()
    assert np.array_equal(y[1], MO.resize_bicubic(big[1], (360, 480), "constant"))


def test_vas_scale_full_suite_vs_oracle(eng):
    """BASELINE config 5: 49x49 model maps scored against 360x480 fixation maps, all six metrics."""
    rng = np.random.RandomState(7)
    B, T, H, W = 2, 8, 360, 480
    gt, _ = SY.make_gazemaps(rng, B, T, 49, 49)
    pred = (gt + 0.3 * rng.uniform(size=gt.shape)).astype(np.float32)
    fix = SY.make_fixationmaps(rng, gt, H, W)
    N = B * T
    pred, gt, fix = pred.reshape(N, 49, 49), gt.reshape(N, 49, 49), fix.reshape(N, H, W)
    fix[3] = 0                                                                 # a frame without fixations
    union = (fix[rng.choice(N, 10, replace=False)] > 0).sum(0).astype(np.float64)
    nfix = fix.reshape(N, -1).sum(1)
    jit = rng.rand(N, H, W)
    ridx = [rng.randint(0, H * W, [int(n), 100]) for n in nfix]
    np.random.seed(11)
    sidx = [EM.shuffled_sample_index(union.ravel() > 0.5, 100, int(n)) if n else np.zeros((0, 100), np.int64) for n in nfix]
    r = eng.evaluate_batch(pred, gt, fix, union=union, jitter_values=jit, rand_index=ridx, shuf_index=sidx)
    for i in range(N):
        kws = {"AUC_Judd": dict(jitter_values=jit[i]), "AUC_Borji": dict(rand_index=ridx[i]),
               "AUC_shuffled": dict(rand_index=sidx[i])}
        for m in EM.ALL_METRICS:
            want = MO.saliency_score_single(m, pred[i].astype(np.float64), gt[i], fix[i], union, **kws.get(m, {}))
            got = r[m][i]
            if np.isnan(want):
                assert np.isnan(got), (m, i)
            else:
                assert abs(got - want) <= (1e-12 if m.startswith("AUC") else 1e-10), (m, i, got, want)
    assert np.isnan(r["AUC_Judd"][3]) and np.isnan(r["NSS"][3]) and not np.isnan(r["sim"][3])


def test_edge_cases(eng):
    rng = np.random.RandomState(2)
    H, W = 30, 41                                     # odd pixel count: scalar (non-vectorised) load path
    pred = rng.uniform(size=(3, H, W)).astype(np.float32)
    gt = rng.uniform(size=(3, H, W)).astype(np.float32)
    fix = np.zeros((3, H, W), np.uint8)
    fix[0, 2, 3] = fix[0, 9, 9] = 1
    fix[2].reshape(-1)[::7] = 1                       # many fixations
    pred[1] = 0.5                                     # constant map: normalize_range -> NaN everywhere
    jit = rng.rand(3, H, W)
    r = eng.evaluate_batch(pred, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter_values=jit)
    for i in (0, 2):
        for m, kw in (("sim", {}), ("cc", {}), ("NSS", {}), ("AUC_Judd", dict(jitter_values=jit[i]))):
            want = MO.saliency_score_single(m, pred[i].astype(np.float64), gt[i], fix[i], **kw)
            assert abs(r[m][i] - want) <= 1e-10, (m, i)
    assert np.isnan(r["sim"][1]) and np.isnan(r["cc"][1]) and np.isnan(r["AUC_Judd"][1])
    # empty batch is a no-op; more fixations than the kernels hold is an error, not a wrong number
    assert eng.to_f64(torch.zeros(0, 4, 4, device="cuda"), True).shape[0] == 0
    with pytest.raises(ValueError):
        eng.evaluate_batch(np.zeros((1, 49, 49), np.float32), np.zeros((1, 49, 49), np.float32),
                           np.ones((1, 49, 49), np.uint8), metrics=("AUC_Judd",))


def test_size_independent_properties_at_scale(eng):
    """N = 512 frames at 360x480 (too slow for the oracle): invariances the metrics must satisfy."""
    rng = np.random.RandomState(5)
    N, H, W = 512, 360, 480
    base, _ = SY.make_gazemaps(rng, 8, 8, 49, 49)
    base = base.reshape(64, 49, 49)
    pred = np.tile(base, (N // 64, 1, 1)) + (0.05 * rng.uniform(size=(N, 49, 49))).astype(np.float32)
    gt = np.tile(base, (N // 64, 1, 1))
    fix = np.tile(SY.make_fixationmaps(rng, base[:, None], H, W).reshape(64, H, W), (N // 64, 1, 1))
    r1 = eng.evaluate_batch(pred, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter=False)
    # affine invariance: the prediction is min-max normalised first, so a*pred+b (a>0) scores the same
    # (a power-of-two scale is exact in fp32 -> identical bits; a general affine map only moves fp32 roundings)
    r2 = eng.evaluate_batch(4.0 * pred, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter=False)
    r2b = eng.evaluate_batch(3.0 * pred + 1.0, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter=False)
    for m in r1:
        assert np.array_equal(r1[m], r2[m]), m
        assert np.allclose(r1[m], r2b[m], rtol=0, atol=1e-3), m
    assert ((r1["AUC_Judd"] >= 0) & (r1["AUC_Judd"] <= 1)).all() and (r1["AUC_Judd"].mean() > 0.8)
    assert ((r1["sim"] > 0) & (r1["sim"] <= 1)).all() and (np.abs(r1["cc"]) <= 1 + 1e-12).all()
    # scoring the ground truth against itself: sim = cc = 1
    r3 = eng.evaluate_batch(gt[:64], gt[:64], fix[:64], metrics=("sim", "cc"))
    assert np.allclose(r3["sim"], 1.0, atol=1e-12) and np.allclose(r3["cc"], 1.0, atol=1e-12)
    # determinism: same inputs, same bits
    r4 = eng.evaluate_batch(pred, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter=False)
    for m in r1:
        assert np.array_equal(r1[m], r4[m]), m


def _vas_case(seed, N, H=360, W=480):
    rng = np.random.RandomState(seed)
    gt, _ = SY.make_gazemaps(rng, 1, N, 49, 49)
    pred = (gt + 0.3 * rng.uniform(size=gt.shape)).astype(np.float32)
    fix = SY.make_fixationmaps(rng, gt, H, W).reshape(N, H, W)
    pred, gt = pred.reshape(N, 49, 49), gt.reshape(N, 49, 49)
    union = (fix[rng.choice(N, min(N, 10), replace=False)] > 0).sum(0).astype(np.float64)
    nfix = fix.reshape(N, -1).sum(1)
    jit = rng.rand(N, H, W)
    ridx = [rng.randint(0, H * W, [int(n), 100]) for n in nfix]
    np.random.seed(seed + 100)
    sidx = [EM.shuffled_sample_index(union.ravel() > 0.5, 100, int(n)) if n else np.zeros((0, 100), np.int64) for n in nfix]
    return pred, gt, fix, union, jit, ridx, sidx


@pytest.mark.parametrize("shape", [(360, 480), (45, 61), (100, 7)])
def test_fused_resize_metrics_equal_two_step_path(shape):
    """rgp_*_resized (enlargement evaluated inside the metric kernels) against rgp_resize_bicubic + the plain kernels:
    thresholds, integer counts and AUCs identical bits; sim / cc / NSS differ only by summation order."""
    H, W = shape
    pred, gt, fix, union, jit, ridx, sidx = _vas_case(3, 12, H, W)
    fix[5] = 0
    a = EM.MetricsEngine(device="cuda", fused="always")
    b = EM.MetricsEngine(device="cuda", fused=False)
    assert a.fused_fits((49, 49), shape)
    kw = dict(union=union, jitter_values=jit, rand_index=ridx, shuf_index=sidx, return_counts=True)
    before = a.lib.launches
    ra = a.evaluate_batch(pred, gt, fix, **kw)
    assert a.lib.launches - before == 1 + 2 + 4        # fixation count, prep x2, fused sim/cc/nss, Judd, Borji, shuffled: no resize launch
    rb = b.evaluate_batch(pred, gt, fix, **kw)
    for k in rb:
        if k in ("sim", "cc", "NSS"):
            assert np.array_equal(np.isnan(ra[k]), np.isnan(rb[k])), k
            m = ~np.isnan(rb[k])
            assert np.abs(ra[k][m] - rb[k][m]).max() <= 1e-13, (k, np.abs(ra[k][m] - rb[k][m]).max())
        else:
            assert np.array_equal(ra[k], rb[k], equal_nan=True), k
    # the (min, max, n_fix) columns of the statistics row are exact as well
    p64 = a.to_f64(torch.from_numpy(pred).cuda(), True)
    g64 = a.to_f64(torch.from_numpy(gt).cuda(), False)
    fx = torch.from_numpy(fix).cuda()
    for mode in ("constant", "nearest"):
        sa = a.sim_cc_nss_resized(p64, g64, fx, shape, mode).cpu().numpy()
        sb = b.sim_cc_nss(b.resize(p64, shape, mode), b.resize(g64, shape, mode), fx).cpu().numpy()
        assert np.array_equal(sa[:, 3:6], sb[:, 3:6])
        assert np.allclose(sa[:, 6:], sb[:, 6:], rtol=1e-13, atol=0)


def test_fused_path_nan_and_large_index_tensor():
    """constant prediction (normalize_range -> NaN map) through the fused kernels; a sampled-AUC index tensor too
    large for the shared-memory stash takes the gather-per-threshold path and gives the same counts."""
    H, W = 120, 160
    pred, gt, fix, union, jit, ridx, sidx = _vas_case(9, 4, H, W)
    pred[1] = 0.25
    a = EM.MetricsEngine(device="cuda", fused="always")
    b = EM.MetricsEngine(device="cuda", fused=False)
    ra = a.evaluate_batch(pred, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter_values=jit)
    rb = b.evaluate_batch(pred, gt, fix, metrics=("sim", "cc", "NSS", "AUC_Judd"), jitter_values=jit)
    for k in ("sim", "cc", "NSS"):
        assert np.isnan(ra[k][1]), k
        assert not np.isnan(ra[k][0]), k
    # AUC-Judd on an all-NaN map: the reference's loop counts 0 pixels above every (NaN) threshold and integrates a
    # finite number (evaluation_metrics.py:84-97); both device paths and the oracle reproduce that number
    assert np.array_equal(ra["AUC_Judd"], rb["AUC_Judd"])
    with np.errstate(invalid="ignore", divide="ignore"):
        want = MO.saliency_score_single("AUC_Judd", pred[1].astype(np.float64), gt[1], fix[1], jitter_values=jit[1])
    assert abs(ra["AUC_Judd"][1] - want) <= 1e-12
    big = [np.random.RandomState(i).randint(0, H * W, [int((fix[i] > 0).sum()), 1500]) for i in range(4)]
    x = a.evaluate_batch(pred, gt, fix, metrics=("AUC_Borji",), rand_index=big, n_rep=1500, return_counts=True)
    y = b.evaluate_batch(pred, gt, fix, metrics=("AUC_Borji",), rand_index=big, n_rep=1500, return_counts=True)
    for k in y:
        assert np.array_equal(x[k], y[k], equal_nan=True), k
    i = 0
    S = MO.normalize_range(MO.resize_bicubic(MO.normalize_range(pred[i].astype(np.float64)), (H, W), "constant")).ravel()
    F = (fix[i] > 0).ravel()
    for rep, (thresholds, tpc, fpc) in enumerate(MO.borji_counts(S[F], S[big[i]], 0.1)):
        assert np.array_equal(x["AUC_Borji_fp"][i, rep, :len(thresholds)], fpc[::-1])


def test_device_rng_kernels_and_fast_evaluation_mode():
    """csrc/rng.cu: the counter-based generator equals its NumPy restatement (tests/emu_lib.py) bit for bit, stays in
    range, and evaluate_batch(rng="device") leaves the deterministic metrics untouched while the sampled AUCs agree with
    the reference-stream path to sampling noise."""
    import ctypes as C
    from emu_lib import EmuLib
    from recurrent_gaze_prediction_b200 import lib as L
    lib = L.get_lib()
    n = 100001
    u = torch.empty(n + 1, dtype=torch.float64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    lib.rgp_fill_uniform_f64(u.data_ptr(), n, 1234567, s)
    ref = np.zeros(n, dtype=np.float64)
    EmuLib().rgp_fill_uniform_f64(ref.ctypes.data, n, 1234567, 0)
    got = u[:n].cpu().numpy()
    assert np.array_equal(got, ref) and got.min() >= 0.0 and got.max() < 1.0 and abs(got.mean() - 0.5) < 5e-3
    nfix = torch.tensor([3, 0, 7, 1], dtype=torch.int32, device="cuda")
    idx = torch.empty(4, 7, 10, dtype=torch.int32, device="cuda")
    lib.rgp_fill_randint_rows(idx.data_ptr(), nfix.data_ptr(), 4, 7, 10, 172800, 99, s)
    ri = np.zeros((4, 7, 10), dtype=np.int32)
    nf = nfix.cpu().numpy()
    EmuLib().rgp_fill_randint_rows(ri.ctypes.data, nf.ctypes.data, 4, 7, 10, 172800, 99, 0)
    g = idx.cpu().numpy()
    assert np.array_equal(g, ri) and (g[1] == -1).all() and (g[0, :3] >= 0).all() and (g[0, 3:] == -1).all() and g.max() < 172800
    # the evaluation call
    rng = np.random.RandomState(3)
    gt, _ = SY.make_gazemaps(rng, 8, 4, 49, 49)
    gt = gt.reshape(32, 49, 49)
    pred = (gt + 0.3 * rng.uniform(size=gt.shape)).astype(np.float32)
    fix = SY.make_fixationmaps(rng, gt[:, None], 120, 160).reshape(32, 120, 160)
    union = (fix[:10].sum(axis=0) > 0).astype(np.uint8)
    eng = EM.MetricsEngine(device="cuda")
    np.random.seed(1)
    a = eng.evaluate_batch(pred, gt, fix, union=union)
    b = eng.evaluate_batch(pred, gt, fix, union=union, rng="device", seed=4)
    for m in ("sim", "cc", "NSS"):
        assert np.array_equal(a[m], b[m], equal_nan=True)
    for m in ("AUC_Judd", "AUC_Borji", "AUC_shuffled"):
        ok = np.isfinite(a[m])
        assert np.abs(a[m][ok] - b[m][ok]).max() < 0.08, m


def _edge_maps(kind, rng, n, h=49, w=49):
    """source maps that drive the lean fused kernels through their other paths (csrc/metrics.cu: zero-minimum passes with
    saturated extrema, the bound-based passes for negative values / a non-zero minimum)."""
    base = rng.uniform(size=(n, h, w))
    if kind == "plateaus":             # exact zeros and an exact plateau at the maximum: clipped-to-range runs, min = 0 attained
        m = np.clip((base - 0.3) / 0.4, 0.0, 1.0)
        m[:, :6] = 0.0
        m[:, -5:, 10:30] = 1.0
        return m, np.clip((rng.uniform(size=(n, h, w)) - 0.5) * 4.0, 0.0, 1.0)
    if kind == "negative":             # values of both signs, non-zero minimum: per-window overshoot bounds
        return base * 3.0 - 0.5, base[::-1] * 2.0 - 0.25
    if kind == "offset":               # strictly positive, minimum > 0
        return base + 0.5, 2.0 + np.sin(base * 7.0)
    raise ValueError(kind)


@pytest.mark.parametrize("kind", ["plateaus", "negative", "offset"])
@pytest.mark.parametrize("shape", [(360, 480), (93, 130), (20, 24)])
def test_lean_fused_kernels_other_ranges_equal_two_step_path(eng, kind, shape):
    """rgp_sim_cc_nss_resized / rgp_auc_judd_resized (lean kernels) against rgp_resize_bicubic + the plain kernels on maps that
    are NOT min-max normalised: negative values, non-zero minima, exact plateaus at both ends of the range; enlarging,
    non-multiple-of-4 widths (byte path of the fixation scan) and shrinking (windows without rows).  Extrema, n_fix,
    thresholds, counts and AUCs bit-identical; sim / cc / NSS within 1e-12 relative (summation order, FMA in sum-only terms)."""
    H, W = shape
    rng = np.random.RandomState(sum(map(ord, kind)) * 1000 + H)
    n = 5
    p, g = _edge_maps(kind, rng, n)
    fix = (rng.uniform(size=(n, H, W)) < 12.0 / (H * W)).astype(np.uint8)
    fix[:, H // 2, W // 3] = 1
    fix[3] = 0                                              # a frame without fixations
    p64 = torch.from_numpy(p).cuda()
    g64 = torch.from_numpy(np.ascontiguousarray(g)).cuda()
    fx = torch.from_numpy(fix).cuda()
    jit = torch.from_numpy(rng.rand(n, H, W)).cuda()
    for mode in ("constant", "nearest"):
        sa = eng.sim_cc_nss_resized(p64, g64, fx, shape, mode).cpu().numpy()
        pr, gr = eng.resize(p64, shape, mode), eng.resize(g64, shape, mode)
        sb = eng.sim_cc_nss(pr, gr, fx).cpu().numpy()
        assert np.array_equal(sa[:, 3:6], sb[:, 3:6]), (mode, sa[:, 3:6], sb[:, 3:6])
        assert np.array_equal(np.isnan(sa), np.isnan(sb))
        ok = ~np.isnan(sb)
        scale = np.maximum(np.abs(sb[ok]), 1.0)
        assert (np.abs(sa[ok] - sb[ok]) / scale).max() <= 1e-12, (mode, np.abs(sa[ok] - sb[ok]).max())
        for jd in (jit, None):
            a1, t1, c1 = eng.auc_judd_resized(p64, fx, shape, jd, 64, mode)
            a2, t2, c2 = eng.auc_judd(pr, fx, jd, 64)
            assert torch.equal(c1, c2) and torch.equal(t1.nan_to_num(-7.0), t2.nan_to_num(-7.0))
            assert np.array_equal(a1.cpu().numpy(), a2.cpu().numpy(), equal_nan=True)


@pytest.mark.parametrize("row", [360 * 480, 49 * 49, 1000003])
def test_count_nonzero_rows_kernel(eng, row):
    """rgp_count_nonzero_rows (fixations per frame) against NumPy: 16-byte path (aligned rows), byte path (odd rows), byte
    values other than 0 / 1."""
    rng = np.random.RandomState(row % 97)
    n = 7
    x = (rng.uniform(size=(n, row)) < 0.01).astype(np.uint8) * rng.randint(1, 256, size=(n, row)).astype(np.uint8)
    x[2] = 0
    x[3, -1] = 255
    d = torch.from_numpy(x).cuda()
    out = torch.full((n,), -5, dtype=torch.int32, device="cuda")
    eng.lib.rgp_count_nonzero_rows(d.data_ptr(), n, row, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert np.array_equal(out.cpu().numpy(), np.count_nonzero(x, axis=1))
