"""CPU oracle (NumPy float64) for the saliency metrics — TEST INFRASTRUCTURE ONLY.

Restates ``/root/reference/evaluation_metrics.py`` for Python 3 / NumPy 2 with
the random draws made explicit (they can be passed in, or are drawn from the
global ``numpy.random`` state in the same order as the reference draws them).
Pinned against the reference module itself by ``oracle/gen_golden_metrics.py``
(the reference file is imported unmodified; see that script for the shims).

Two things are NOT in the reference and are defined here (parity unpinned):
* ``resize_bicubic`` — stands in for ``skimage.transform.resize(order=3)``
  (scikit-image is un-vendored and its version unpinned): half-pixel-centre
  coordinate map, Catmull-Rom (Keys a=-0.5) 4x4 taps, boundary 'constant'
  (cval 0) or 'nearest', result clipped to the input range (skimage clip=True).
  The operation order is spelled out so the CUDA kernel can match it bit-for-bit.
* ``nss`` — the reference has no NSS; definition follows the upstream the file
  credits (evaluation_metrics.py:5): z-score the saliency map (population std),
  mean at fixated pixels.
"""
from __future__ import annotations

import numpy as np
import numpy.random as random


def normalize_range(x):
    """evaluation_metrics.py:15-17 (NaN on constant maps, by design)."""
    x = np.asarray(x, dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        return (x - np.min(x)) / (np.max(x) - np.min(x))


def resize_onehot_tensor_sparse(x, target_shape):
    """evaluation_metrics.py:19-39."""
    assert len(target_shape) == 2
    H1, W1 = x.shape[-2:]
    H2, W2 = target_shape
    if x.ndim != 2:
        raise ValueError("x.shape : %s" % (x.shape,))
    ret = np.zeros((H2, W2), dtype=bool)
    for y, xx in zip(*np.where(x > 0)):
        y_ = y * (H2 - 1.0) / (H1 - 1.0)
        x_ = xx * (W2 - 1.0) / (W1 - 1.0)
        ret[int(np.round(y_) + 1e-9), int(np.round(x_) + 1e-9)] = 1
    return ret


# ----------------------------------------------------------------------------
# resize (stand-in for skimage.transform.resize(order=3)); see module docstring
# ----------------------------------------------------------------------------
def _cubic_weights(t):
    """Catmull-Rom weights for taps at -1,0,1,2 around floor(coord); fixed op order."""
    t2 = t * t
    t3 = t2 * t
    w0 = 0.5 * (-t3 + 2.0 * t2 - t)
    w1 = 0.5 * (3.0 * t3 - 5.0 * t2 + 2.0)
    w2 = 0.5 * (-3.0 * t3 + 4.0 * t2 + t)
    w3 = 0.5 * (t3 - t2)
    return w0, w1, w2, w3


def resize_bicubic(img, out_shape, mode="constant"):
    """out[r,c] = sum_i wr_i * (sum_j wc_j * src[r0+i, c0+j]); rows combined after columns,
    taps accumulated left-to-right; coordinate = (o + 0.5) * (in/out) - 0.5."""
    img = np.asarray(img, dtype=np.float64)
    H, W = img.shape
    OH, OW = out_shape
    if (H, W) == (OH, OW):
        return img.copy()
    rs = (np.arange(OH, dtype=np.float64) + 0.5) * (float(H) / OH) - 0.5
    cs = (np.arange(OW, dtype=np.float64) + 0.5) * (float(W) / OW) - 0.5
    r0 = np.floor(rs)
    c0 = np.floor(cs)
    wr = _cubic_weights(rs - r0)
    wc = _cubic_weights(cs - c0)
    r0 = r0.astype(np.int64) - 1
    c0 = c0.astype(np.int64) - 1

    def fetch(ri, ci):
        if mode == "nearest":
            return img[np.clip(ri, 0, H - 1)[:, None], np.clip(ci, 0, W - 1)[None, :]]
        ok = ((ri >= 0) & (ri < H))[:, None] & ((ci >= 0) & (ci < W))[None, :]
        v = img[np.clip(ri, 0, H - 1)[:, None], np.clip(ci, 0, W - 1)[None, :]]
        return np.where(ok, v, 0.0)

    out = np.zeros((OH, OW), dtype=np.float64)
    for i in range(4):
        row = np.zeros((OH, OW), dtype=np.float64)
        for j in range(4):
            row = row + wc[j][None, :] * fetch(r0 + i, c0 + j)
        out = out + wr[i][:, None] * row
    lo, hi = img.min(), img.max()
    return np.minimum(np.maximum(out, lo), hi)


# ----------------------------------------------------------------------------
# AUC family
# ----------------------------------------------------------------------------
def auc_judd_counts(S, F):
    """Integer core of AUC_Judd (evaluation_metrics.py:83-97): for fixation values sorted
    descending, above_th[k] = #{S >= thr_k}.  Returned for the bit-exact parity check."""
    S_fix = S[F]
    thresholds = np.sort(S_fix)[::-1]
    above = np.array([int(np.sum(S >= t)) for t in thresholds], dtype=np.int64)
    return thresholds, above


def AUC_Judd(fixation_map, saliency_map, jitter=True, jitter_values=None, return_counts=False):
    """evaluation_metrics.py:42-98. `jitter_values` = the random.rand(H,W) draw (before *1e-7);
    drawn from the global RNG when None, exactly where the reference draws it."""
    saliency_map = np.array(saliency_map, dtype=np.float64, copy=True)
    fixation_map = np.asarray(fixation_map) > 0.5
    if not np.any(fixation_map):
        return (np.nan, None) if return_counts else np.nan
    if saliency_map.shape != fixation_map.shape:
        saliency_map = resize_bicubic(saliency_map, fixation_map.shape, mode="nearest")
    if jitter:
        if jitter_values is None:
            jitter_values = random.rand(*saliency_map.shape)
        saliency_map = saliency_map + jitter_values * 1e-7
    saliency_map = normalize_range(saliency_map)
    S = saliency_map.ravel()
    F = fixation_map.ravel()
    n_fix = int(F.sum())
    n_pixels = len(S)
    thresholds, above = auc_judd_counts(S, F)
    tp = np.zeros(n_fix + 2)
    fp = np.zeros(n_fix + 2)
    tp[-1] = 1
    fp[-1] = 1
    for k in range(n_fix):
        tp[k + 1] = (k + 1) / float(n_fix)
        fp[k + 1] = (above[k] - k - 1) / float(n_pixels - n_fix)
    auc = _trapz(tp, fp)
    return (auc, above) if return_counts else auc


def _trapz(y, x):
    """np.trapz(y, x) (removed from NumPy 2's namespace as trapz; same arithmetic)."""
    y = np.asarray(y, dtype=np.float64)
    x = np.asarray(x, dtype=np.float64)
    d = np.diff(x)
    return float((d * (y[1:] + y[:-1]) / 2.0).sum())


def borji_counts(S_fix, S_rand, step_size=0.1):
    """Integer core of AUC_Borji (evaluation_metrics.py:152-163): per repetition, the
    threshold list and the tp/fp integer counts."""
    n_rep = S_rand.shape[1]
    out = []
    for rep in range(n_rep):
        mx = np.max(np.r_[S_fix, S_rand[:, rep]])
        thresholds = np.arange(0, mx, step_size)[::-1]
        tpc = np.array([int(np.sum(S_fix >= t)) for t in thresholds], dtype=np.int64)
        fpc = np.array([int(np.sum(S_rand[:, rep] >= t)) for t in thresholds], dtype=np.int64)
        out.append((thresholds, tpc, fpc))
    return out


def AUC_Borji(fixation_map, saliency_map, n_rep=100, step_size=0.1, rand_sampler=None, rand_index=None):
    """evaluation_metrics.py:101-164. `rand_index` [n_fix, n_rep] = the random.randint draw
    (global RNG when None). `rand_sampler(S, F, n_rep, n_fix)` as in the reference."""
    saliency_map = np.array(saliency_map, dtype=np.float64, copy=True)
    fixation_map = np.asarray(fixation_map) > 0.5
    if not np.any(fixation_map):
        return np.nan
    if saliency_map.shape != fixation_map.shape:
        saliency_map = resize_bicubic(saliency_map, fixation_map.shape, mode="nearest")
    saliency_map = normalize_range(saliency_map)
    S = saliency_map.ravel()
    F = fixation_map.ravel()
    S_fix = S[F]
    n_fix = len(S_fix)
    n_pixels = len(S)
    if rand_sampler is None:
        if rand_index is None:
            rand_index = random.randint(0, n_pixels, [n_fix, n_rep])
        S_rand = S[rand_index]
    else:
        S_rand = rand_sampler(S, F, n_rep, n_fix)
    auc = np.zeros(n_rep) * np.nan
    for rep, (thresholds, tpc, fpc) in enumerate(borji_counts(S_fix, S_rand, step_size)):
        tp = np.zeros(len(thresholds) + 2)
        fp = np.zeros(len(thresholds) + 2)
        tp[-1] = 1
        fp[-1] = 1
        tp[1:-1] = tpc / float(n_fix)
        fp[1:-1] = fpc / float(n_fix)
        auc[rep] = _trapz(tp, fp)
    return float(np.mean(auc))


def shuffled_sample_index(other_flat, n_rep, n_fix):
    """sample_other of AUC_shuffled (evaluation_metrics.py:197-202): per repetition
    random.permutation(len(fixated))[:n_fix], transposed to [<=n_fix, n_rep] pixel indices.
    (The reference's py2 `map` over np.tile(range(n), [n_rep,1]) permutes each row.)"""
    fixated = np.nonzero(other_flat)[0]
    indexer = [random.permutation(len(fixated))[:n_fix] for _ in range(n_rep)]
    return fixated[np.transpose(indexer)]


def AUC_shuffled(fixation_map, saliency_map, other_map, n_rep=100, step_size=0.1, rand_index=None):
    """evaluation_metrics.py:167-204."""
    other_map = np.asarray(other_map) > 0.5
    if other_map.shape != np.asarray(fixation_map).shape:
        raise ValueError("other_map.shape != fixation_map.shape")

    def sample_other(S, F, n_rep_, n_fix):
        r = rand_index if rand_index is not None else shuffled_sample_index(other_map.ravel(), n_rep_, n_fix)
        return S[r]

    return AUC_Borji(fixation_map, saliency_map, n_rep, step_size, sample_other)


# ----------------------------------------------------------------------------
# distribution metrics
# ----------------------------------------------------------------------------
def similarity(gtsAnn, resAnn):
    """evaluation_metrics.py:207-218."""
    gtsAnn = np.asarray(gtsAnn, dtype=np.float64)
    resAnn = np.asarray(resAnn, dtype=np.float64)
    return float(np.minimum(gtsAnn / gtsAnn.sum(), resAnn / resAnn.sum()).sum())


def cc(gtsAnn, resAnn):
    """evaluation_metrics.py:221-236 (z-score with population std, then np.corrcoef)."""
    gtsAnn = np.asarray(gtsAnn, dtype=np.float64)
    resAnn = np.asarray(resAnn, dtype=np.float64)
    fixationMap = gtsAnn - np.mean(gtsAnn)
    if np.max(fixationMap) > 0:
        fixationMap = fixationMap / np.std(fixationMap)
    salMap = resAnn - np.mean(resAnn)
    if np.max(salMap) > 0:
        salMap = salMap / np.std(salMap)
    return float(np.corrcoef(salMap.reshape(-1), fixationMap.reshape(-1))[0][1])


def nss(fixation_map, saliency_map):
    """NOT in the reference (parity unpinned): mean of the z-scored saliency map at fixations."""
    S = np.asarray(saliency_map, dtype=np.float64)
    F = np.asarray(fixation_map) > 0.5
    if not np.any(F):
        return np.nan
    if S.shape != F.shape:
        S = resize_bicubic(S, F.shape, mode="nearest")
    z = (S - S.mean()) / S.std()
    return float(z[F].mean())


def saliency_score_single(metric, pred_map, gt_map, fixation_map, other_map_union=None, **rand):
    """evaluation_metrics.py:239-272."""
    fixation_map = np.asarray(fixation_map)
    pred_map = normalize_range(pred_map)
    pred_map_orig = resize_bicubic(pred_map, fixation_map.shape, mode="constant")
    gt_map_orig = resize_bicubic(np.asarray(gt_map, dtype=np.float64), fixation_map.shape, mode="constant")
    if metric == "cc":
        return cc(gt_map_orig, pred_map_orig)
    if metric == "sim":
        return similarity(gt_map_orig, pred_map_orig)
    if metric == "AUC_Judd":
        return AUC_Judd(fixation_map, pred_map_orig, jitter_values=rand.get("jitter_values"))
    if metric == "AUC_Borji":
        return AUC_Borji(fixation_map, pred_map_orig, rand_index=rand.get("rand_index"))
    if metric == "AUC_shuffled":
        if other_map_union is None:
            raise ValueError("other_map_union required")
        return AUC_shuffled(fixation_map, pred_map_orig, other_map_union, rand_index=rand.get("rand_index"))
    if metric == "NSS":
        return nss(fixation_map, pred_map_orig)
    raise ValueError(metric)


def other_map_union(fixation_maps, M=10):
    """evaluation_metrics.py:280-287."""
    assert len(fixation_maps) >= M
    u = np.zeros(np.asarray(fixation_maps[0]).shape)
    for i in random.choice(range(len(fixation_maps)), M, replace=False):
        u += (np.asarray(fixation_maps[i]) > 0).astype(np.int64)
    return u


def saliency_score(metric, pred_maps, gt_maps, fixation_maps):
    """evaluation_metrics.py:275-295 (NaN-propagating mean)."""
    assert len(gt_maps) == len(pred_maps) == len(fixation_maps)
    union = other_map_union(fixation_maps)
    scores = [saliency_score_single(metric, p, g, f, union) for g, p, f in zip(gt_maps, pred_maps, fixation_maps)]
    return float(np.mean(scores))


AVAILABLE_METRICS = ("sim", "cc", "AUC_shuffled", "AUC_Borji")
