"""Saliency evaluation metrics on the GPU, behind the reference's ``evaluation_metrics`` interface.

Mirrors ``/root/reference/evaluation_metrics.py`` (names, argument meaning, NaN / ValueError behaviour):
``saliency_score`` (:275-295), ``saliency_score_single`` (:239-272), ``AUC_Judd`` (:42-98), ``AUC_Borji``
(:101-164), ``AUC_shuffled`` (:167-204), ``similarity`` (:207-218), ``cc`` (:221-236),
``resize_onehot_tensor_sparse`` (:19-39), ``normalize_range`` (:15-17), ``AVAILABLE_METRICS`` (:297), plus
``nss`` (the north star asks for it; the reference has none — definition in DESIGN.md).

All map arithmetic runs in the float64 kernels of ``csrc/metrics.cu`` through the C-ABI; there is no CPU
fallback.  What stays on the host is what the reference also does on the host *outside* the map arithmetic:
drawing the random jitter / negative-sample indices from the **global** ``numpy.random`` state, in the
reference's order, so that a seeded run draws the same numbers; the drawn arrays are shipped to the device.

``MetricsEngine.evaluate_batch`` is the fast path (N frames per launch); the reference-style functions are thin
single-frame wrappers over it.
"""
from __future__ import annotations

import numpy as np
import numpy.random as random
import torch

from . import lib as L

AVAILABLE_METRICS = ['sim', 'cc', 'AUC_shuffled', 'AUC_Borji']          # evaluation_metrics.py:297
ALL_METRICS = ('sim', 'cc', 'NSS', 'AUC_Judd', 'AUC_Borji', 'AUC_shuffled')
KMAXFIX = 1024                                                          # csrc/metrics.cu kMaxFix


def normalize_range(x):
    """evaluation_metrics.py:15-17 (NaN on a constant map, by design)."""
    x = np.asarray(x, dtype=np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        return (x - np.min(x)) / (np.max(x) - np.min(x))


def resize_onehot_tensor_sparse(x, target_shape):
    """evaluation_metrics.py:19-39 — host index arithmetic on the (few) fixated pixels."""
    assert len(target_shape) == 2
    x = np.asarray(x)
    if x.ndim != 2:
        raise ValueError("x.shape : %s" % (x.shape,))
    H1, W1 = x.shape
    H2, W2 = target_shape
    ret = np.zeros((H2, W2), dtype=bool)
    ys, xs = np.where(x > 0)
    yy = (np.round(ys * (H2 - 1.0) / (H1 - 1.0)) + 1e-9).astype(np.int64)
    xx = (np.round(xs * (W2 - 1.0) / (W1 - 1.0)) + 1e-9).astype(np.int64)
    ret[yy, xx] = True
    return ret


def shuffled_sample_index(other_flat, n_rep, n_fix):
    """The draw of AUC_shuffled's sample_other (evaluation_metrics.py:197-202): per repetition
    ``random.permutation(len(fixated))[:n_fix]`` -> pixel indices [<=n_fix, n_rep]."""
    fixated = np.nonzero(np.asarray(other_flat).ravel())[0]
    indexer = [random.permutation(len(fixated))[:n_fix] for _ in range(n_rep)]
    return fixated[np.transpose(indexer)]


class MetricsEngine(object):
    """Batched device evaluation of the saliency metrics for N frames at once."""

    def __init__(self, device="cuda", lib=None, chunk=256, fused=True):
        self.device = torch.device(device)
        self.lib = lib if lib is not None else L.get_lib()
        self.chunk = chunk
        # fused: the bicubic enlargement of the model maps is evaluated inside the metric kernels (rgp_*_resized) and
        # never written to HBM.  True / "always" = whenever the shape fits the kernels' tables: with the lean kernels of
        # round 2 that is the faster path for every metric set (profiles/r02_notes.md, passes 25-33; in round 1 the
        # two-step path won when AUC_Judd was requested).  False forces rgp_resize_bicubic + the plain kernels.
        self.fused = fused

    def _s(self):
        if self.device.type == "cuda":
            return torch.cuda.current_stream(self.device).cuda_stream
        return 0

    def _dev(self, x, dtype):
        if isinstance(x, torch.Tensor):
            return x.to(device=self.device, dtype=dtype).contiguous()
        return torch.from_numpy(np.ascontiguousarray(x)).to(device=self.device, dtype=dtype)

    # ------------------------------------------------------------------ device stages
    def to_f64(self, maps_f32, normalize):
        """fp32 [n,h,w] -> float64, optionally min-max normalised per frame (evaluation_metrics.py:246)."""
        n, h, w = maps_f32.shape
        out = torch.empty(n, h, w, dtype=torch.float64, device=self.device)
        self.lib.rgp_metrics_prep(maps_f32.data_ptr(), out.data_ptr(), n, h * w, 1 if normalize else 0, self._s())
        return out

    def resize(self, maps_f64, out_shape, mode="constant"):
        """evaluation_metrics.py:248,255 (`mode='constant'`) and :76,137 (`mode='nearest'`)."""
        n, h, w = maps_f64.shape
        oh, ow = out_shape
        out = torch.empty(n, oh, ow, dtype=torch.float64, device=self.device)
        self.lib.rgp_resize_bicubic(maps_f64.data_ptr(), out.data_ptr(), n, h, w, oh, ow,
                                    1 if mode == "nearest" else 0, self._s())
        return out

    def sim_cc_nss(self, pred_f64, gt_f64, fix_u8):
        n = pred_f64.shape[0]
        hw = pred_f64[0].numel()
        out = torch.empty(n, 8, dtype=torch.float64, device=self.device)
        self.lib.rgp_sim_cc_nss(pred_f64.data_ptr(), gt_f64.data_ptr(), fix_u8.data_ptr() if fix_u8 is not None else 0,
                                out.data_ptr(), n, hw, self._s())
        return out

    def auc_judd(self, sal_f64, fix_u8, jitter_f64=None, maxfix=64):
        n = sal_f64.shape[0]
        hw = sal_f64[0].numel()
        thr = torch.empty(n, maxfix, dtype=torch.float64, device=self.device)
        counts = torch.empty(n, maxfix, dtype=torch.int64, device=self.device)
        auc = torch.empty(n, dtype=torch.float64, device=self.device)
        self.lib.rgp_auc_judd(sal_f64.data_ptr(), jitter_f64.data_ptr() if jitter_f64 is not None else 0,
                              fix_u8.data_ptr(), thr.data_ptr(), counts.data_ptr(), auc.data_ptr(), n, hw, maxfix,
                              self._s())
        return auc, thr, counts

    def auc_sampled(self, sal_f64, fix_u8, idx_i32, minmax, mm_stride, step=0.1, maxfix=64, want_counts=False):
        n = sal_f64.shape[0]
        hw = sal_f64[0].numel()
        _, nrows, nrep = idx_i32.shape
        nthr = int(np.ceil(1.0 / step)) + 1
        auc = torch.empty(n, dtype=torch.float64, device=self.device)
        tpc = fpc = None
        if want_counts:
            tpc = torch.empty(n, nrep, nthr, dtype=torch.int32, device=self.device)
            fpc = torch.empty(n, nrep, nthr, dtype=torch.int32, device=self.device)
        self.lib.rgp_auc_sampled(sal_f64.data_ptr(), fix_u8.data_ptr(), idx_i32.data_ptr(), minmax.data_ptr(),
                                 mm_stride, auc.data_ptr(), tpc.data_ptr() if want_counts else 0,
                                 fpc.data_ptr() if want_counts else 0, n, hw, nrows, nrep, float(step), nthr,
                                 maxfix, self._s())
        return auc, tpc, fpc

    # ------------------------------------------------------------------ fused resize + metric (csrc/metrics.cu)
    def fused_fits(self, src_shape, out_shape):
        return bool(self.lib.rgp_metrics_fused_fits(int(src_shape[0]), int(src_shape[1]), int(out_shape[0]),
                                                    int(out_shape[1])))

    def sim_cc_nss_resized(self, pred_src, gt_src, fix_u8, out_shape, mode="constant"):
        n, h, w = pred_src.shape
        oh, ow = out_shape
        out = torch.empty(n, 8, dtype=torch.float64, device=self.device)
        self.lib.rgp_sim_cc_nss_resized(pred_src.data_ptr(), gt_src.data_ptr(),
                                        fix_u8.data_ptr() if fix_u8 is not None else 0, out.data_ptr(), n, h, w, oh, ow,
                                        1 if mode == "nearest" else 0, self._s())
        return out

    def auc_judd_resized(self, sal_src, fix_u8, out_shape, jitter_f64=None, maxfix=64, mode="constant"):
        n, h, w = sal_src.shape
        oh, ow = out_shape
        thr = torch.empty(n, maxfix, dtype=torch.float64, device=self.device)
        counts = torch.empty(n, maxfix, dtype=torch.int64, device=self.device)
        auc = torch.empty(n, dtype=torch.float64, device=self.device)
        self.lib.rgp_auc_judd_resized(sal_src.data_ptr(), jitter_f64.data_ptr() if jitter_f64 is not None else 0,
                                      fix_u8.data_ptr(), thr.data_ptr(), counts.data_ptr(), auc.data_ptr(), n, h, w,
                                      oh, ow, 1 if mode == "nearest" else 0, maxfix, self._s())
        return auc, thr, counts

    def auc_sampled_resized(self, sal_src, fix_u8, idx_i32, minmax, mm_stride, out_shape, step=0.1, maxfix=64,
                            want_counts=False, mode="constant"):
        n, h, w = sal_src.shape
        oh, ow = out_shape
        _, nrows, nrep = idx_i32.shape
        nthr = int(np.ceil(1.0 / step)) + 1
        auc = torch.empty(n, dtype=torch.float64, device=self.device)
        tpc = fpc = None
        if want_counts:
            tpc = torch.empty(n, nrep, nthr, dtype=torch.int32, device=self.device)
            fpc = torch.empty(n, nrep, nthr, dtype=torch.int32, device=self.device)
        self.lib.rgp_auc_sampled_resized(sal_src.data_ptr(), fix_u8.data_ptr(), idx_i32.data_ptr(), minmax.data_ptr(),
                                         mm_stride, auc.data_ptr(), tpc.data_ptr() if want_counts else 0,
                                         fpc.data_ptr() if want_counts else 0, n, h, w, oh, ow,
                                         1 if mode == "nearest" else 0, nrows, nrep, float(step), nthr, maxfix,
                                         self._s())
        return auc, tpc, fpc

    # ------------------------------------------------------------------ host draws (global numpy RNG, reference order)
    @staticmethod
    def _pad_index(rows_list, nrep):
        """list of [n_i, nrep] int arrays -> int32 [N, max n_i, nrep], -1 padded"""
        nmax = max([r.shape[0] for r in rows_list] + [1])
        out = np.full((len(rows_list), nmax, nrep), -1, dtype=np.int32)
        for i, r in enumerate(rows_list):
            if r.size:
                out[i, :r.shape[0]] = r
        return out

    def evaluate_batch(self, pred, gt, fix, union=None, metrics=ALL_METRICS, n_rep=100, step=0.1,
                       jitter=True, jitter_values=None, rand_index=None, shuf_index=None, pred_is_normalized=False,
                       return_counts=False, rng="numpy", seed=0):
        """pred, gt: [N,h,w] fp32 (numpy or device tensor) model-resolution maps; fix: [N,H,W] uint8 / bool;
        union: [H,W] other-fixation union (AUC_shuffled).  Follows saliency_score_single
        (evaluation_metrics.py:239-272) per frame.  Random draws: pass `jitter_values` [N,H,W] / `rand_index`
        (list of [n_fix_i,n_rep]) / `shuf_index` (list of [<=n_fix_i,n_rep]) or let them be drawn from the global
        numpy RNG, frame by frame, in the order Borji -> Judd -> shuffled (evaluate_gaze.py:123-131).
        rng="device" (opt-in, NOT the reference's stream): the jitter and the Borji negatives are generated on the device
        by a counter-based hash of `seed` (rgp_fill_uniform_f64 / rgp_fill_randint_rows) and the shuffled-AUC picks by a
        vectorised NumPy Generator - statistically equivalent draws without the 1.4 GB of host doubles per 1024 frames
        that reproducing NumPy's global stream costs (99 % of this call's wall time at 360x480).
        Returns {metric: float64 ndarray [N]} (+ '<metric>_counts' with return_counts)."""
        if rng not in ("numpy", "device"):
            raise ValueError("rng must be 'numpy' (the reference's global stream) or 'device'")
        dev_rng = rng == "device"
        for m in metrics:
            if m not in ALL_METRICS:
                raise ValueError("Unknown metric : %s" % m)
        fix_h = _dense(fix.cpu() if isinstance(fix, torch.Tensor) else fix)
        # one byte per pixel, non-zero = fixated (`> 0.5` of the reference, evaluation_metrics.py:241; a bool / uint8 array is
        # taken as it is: the two extra passes over 177 MB per 1024 frames at 360x480 were a third of the device-RNG call)
        if fix_h.dtype == np.bool_:
            fix_h = fix_h.view(np.uint8)
        elif fix_h.dtype != np.uint8:
            fix_h = (fix_h > 0.5).view(np.uint8)
        N, H, W = fix_h.shape
        if len(pred) != N or len(gt) != N:
            raise ValueError("pred/gt/fixation maps differ in length")
        # fixations per frame, counted on the device from the bytes the metric kernels read anyway (the fixation maps go up
        # once, 177 MB per 1024 frames at 360x480; NumPy's per-row count of the same array costs 5x the upload)
        fix_dev = self._dev(fix_h, torch.uint8)
        nfix_i32 = torch.empty(N, dtype=torch.int32, device=self.device)
        self.lib.rgp_count_nonzero_rows(fix_dev.data_ptr(), N, H * W, nfix_i32.data_ptr(), self._s())
        nfix = nfix_i32.cpu().numpy().astype(np.int64)
        if nfix.max(initial=0) > KMAXFIX:
            raise ValueError("a frame has %d fixations; the AUC kernels hold at most %d" % (nfix.max(), KMAXFIX))
        maxfix = int(max(1, nfix.max(initial=0)))
        need_shuf = 'AUC_shuffled' in metrics
        if need_shuf:
            if union is None:
                raise ValueError("other_map_union required for AUC_shuffled")
            union_h = np.asarray(union) > 0.5
            if union_h.shape != (H, W):
                raise ValueError("other_map.shape != fixation_map.shape")
        # ---- host draws, frame by frame, reference order
        npx = H * W
        draw_b = 'AUC_Borji' in metrics and rand_index is None and not dev_rng
        draw_j = 'AUC_Judd' in metrics and jitter and jitter_values is None and not dev_rng
        draw_s = need_shuf and shuf_index is None and not dev_rng
        dev_b = dev_rng and 'AUC_Borji' in metrics and rand_index is None
        dev_j = dev_rng and 'AUC_Judd' in metrics and jitter and jitter_values is None
        if dev_rng and need_shuf and shuf_index is None:
            # n_fix distinct picks of the union's fixated pixels per repetition = the n_fix smallest of one random key per
            # union pixel (a uniformly random subset).  Vectorised over frames: one partition per (frame, repetition) row
            # at the largest n_fix, then a sort of that short prefix so that the first k of it are the k smallest keys.
            gen = np.random.Generator(np.random.PCG64(int(seed) + 0x5EED))
            fixated = np.nonzero(union_h.ravel())[0]
            L_ = len(fixated)
            kk = np.minimum(nfix, L_).astype(np.int64)
            kmax = int(kk.max(initial=0))
            shuf_index = [np.zeros((0, n_rep), dtype=np.int64)] * N
            if kmax > 0:
                per = max(1, int(64e6 // max(1, n_rep * L_ * 4)))            # frames per block: <= 64 MB of keys
                for b0 in range(0, N, per):
                    b1 = min(N, b0 + per)
                    keys = gen.random((b1 - b0, n_rep, L_), dtype=np.float32)
                    if kmax < L_:
                        part = np.argpartition(keys, kmax - 1, axis=2)[:, :, :kmax]
                    else:
                        part = np.broadcast_to(np.arange(L_), keys.shape).copy()
                    pk = np.take_along_axis(keys, part, axis=2)
                    part = np.take_along_axis(part, np.argsort(pk, axis=2), axis=2)      # [frames, n_rep, kmax]
                    for i in range(b0, b1):
                        if kk[i] > 0:
                            shuf_index[i] = fixated[part[i - b0, :, :kk[i]].T]
        nfix_dev = nfix_i32 if dev_b else None
        if draw_b:
            rand_index = []
        if draw_j:
            jitter_values = np.zeros((N, H, W), dtype=np.float64)
        if draw_s:
            shuf_index = []
        if draw_b or draw_j or draw_s:
            for i in range(N):
                has = nfix[i] > 0          # the reference returns NaN before drawing when a frame has no fixation
                if draw_b:
                    rand_index.append(random.randint(0, npx, [int(nfix[i]), n_rep]) if has
                                      else np.zeros((0, n_rep), dtype=np.int64))
                if draw_j and has:
                    jitter_values[i] = random.rand(H, W)
                if draw_s:
                    shuf_index.append(shuffled_sample_index(union_h.ravel(), n_rep, int(nfix[i])) if has
                                      else np.zeros((0, n_rep), dtype=np.int64))
        out = {m: np.empty(N, dtype=np.float64) for m in metrics}
        extra = {}
        for c0 in range(0, N, self.chunk):
            c1 = min(N, c0 + self.chunk)
            sl = slice(c0, c1)
            p32 = self._dev(pred[sl], torch.float32)
            g32 = self._dev(gt[sl], torch.float32)
            fx = fix_dev[sl]
            p64 = self.to_f64(p32, normalize=not pred_is_normalized)
            g64 = self.to_f64(g32, normalize=False)
            # maps below fixation resolution: enlargement fused into the metric kernels (never materialised)
            # maps below fixation resolution: the enlargement is evaluated inside the metric kernels (never materialised).
            # With the lean kernels of round 2 this is the faster path for every metric set, AUC-Judd included
            # (DESIGN.md section 5); fused=False forces the two-step path.
            fused = bool(self.fused) and tuple(p64.shape[1:]) != (H, W) and self.fused_fits(p64.shape[1:], (H, W))
            if fused:
                stats = self.sim_cc_nss_resized(p64, g64, fx, (H, W))
            else:
                pr = self.resize(p64, (H, W), "constant")
                gr = self.resize(g64, (H, W), "constant")
                stats = self.sim_cc_nss(pr, gr, fx)
            res = {}
            if 'AUC_Judd' in metrics:
                if dev_j:
                    jd = torch.empty(c1 - c0, H, W, dtype=torch.float64, device=self.device)
                    self.lib.rgp_fill_uniform_f64(jd.data_ptr(), jd.numel(), (int(seed) * 1000003 + c0) & (2 ** 63 - 1), self._s())
                else:
                    jd = self._dev(jitter_values[sl], torch.float64) if (jitter and jitter_values is not None) else None
                if fused:
                    res['AUC_Judd'], thr, cnt = self.auc_judd_resized(p64, fx, (H, W), jd, maxfix)
                else:
                    res['AUC_Judd'], thr, cnt = self.auc_judd(pr, fx, jd, maxfix)
                if return_counts:
                    extra.setdefault('AUC_Judd_counts', []).append(cnt.cpu().numpy())
                    extra.setdefault('AUC_Judd_thresholds', []).append(thr.cpu().numpy())
            for name, index in (('AUC_Borji', rand_index), ('AUC_shuffled', shuf_index)):
                if name in metrics:
                    if name == 'AUC_Borji' and dev_b:
                        idx = torch.empty(c1 - c0, maxfix, n_rep, dtype=torch.int32, device=self.device)
                        self.lib.rgp_fill_randint_rows(idx.data_ptr(), nfix_dev[c0:c1].contiguous().data_ptr(), c1 - c0, maxfix,
                                                       n_rep, npx, (int(seed) * 7919 + 17 + c0) & (2 ** 63 - 1), self._s())
                    else:
                        idx = self._dev(self._pad_index([np.asarray(r) for r in index[c0:c1]], n_rep), torch.int32)
                    mm = stats[:, 3:]
                    if fused:
                        auc, tpc, fpc = self.auc_sampled_resized(p64, fx, idx, mm, 8, (H, W), step, maxfix,
                                                                 want_counts=return_counts)
                    else:
                        auc, tpc, fpc = self.auc_sampled(pr, fx, idx, mm, 8, step, maxfix, want_counts=return_counts)
                    res[name] = auc
                    if return_counts:
                        extra.setdefault(name + '_tp', []).append(tpc.cpu().numpy())
                        extra.setdefault(name + '_fp', []).append(fpc.cpu().numpy())
            st = stats.cpu().numpy()
            for m in metrics:
                if m == 'sim':
                    out[m][sl] = st[:, 0]
                elif m == 'cc':
                    out[m][sl] = st[:, 1]
                elif m == 'NSS':
                    out[m][sl] = st[:, 2]
                else:
                    a = res[m].cpu().numpy()
                    if (a == -1.0).any():
                        raise ValueError("%s: more fixations than maxfix=%d" % (m, maxfix))
                    out[m][sl] = a
        for k, v in extra.items():
            out[k] = np.concatenate(v, axis=0)
        return out


_ENGINE = None


def default_engine():
    global _ENGINE
    if _ENGINE is None:
        _ENGINE = MetricsEngine()
    return _ENGINE


def _as_maps(x):
    return np.asarray(x, dtype=np.float32)[None]


def _dense(x):
    """evaluation_metrics.py:240: `if scipy.sparse.issparse(fixation_map): fixation_map = fixation_map.toarray()` — the
    reference's loaders emit sparse fixation maps (crc_input_data_seq.py:277-321).  Also accepts a list of them."""
    if hasattr(x, "toarray"):
        return x.toarray()
    if isinstance(x, (list, tuple)) and len(x) and hasattr(x[0], "toarray"):
        return np.stack([f.toarray() for f in x])
    return np.asarray(x)


def _one(metric, fixation_map, saliency_map, gt_map=None, **kw):
    """single pre-normalised-or-not frame through the batch path with `saliency_map` used as is
    (the standalone AUC_* / similarity / cc entry points take maps that are already prepared)."""
    eng = kw.pop("engine", None) or default_engine()
    fixation_map = _dense(fixation_map)
    sal = np.asarray(saliency_map, dtype=np.float64)
    F = fixation_map > 0.5
    dev = eng.device
    s = torch.from_numpy(np.ascontiguousarray(sal))[None].to(dev)
    if sal.shape != F.shape:     # evaluation_metrics.py:74-76,135-137
        s = eng.resize(s, F.shape, "nearest")
    fx = torch.from_numpy(F.astype(np.uint8))[None].to(dev)
    return eng, s, fx, F


def AUC_Judd(fixation_map, saliency_map, jitter=True, jitter_values=None, engine=None):
    """evaluation_metrics.py:42-98.  NaN (and a message) when the frame has no fixation."""
    eng, s, fx, F = _one('AUC_Judd', fixation_map, saliency_map, engine=engine)
    n_fix = int(F.sum())
    if n_fix == 0:
        print('no fixation to predict')
        return np.nan
    if n_fix > KMAXFIX:
        raise ValueError("more than %d fixations" % KMAXFIX)
    jd = None
    if jitter:
        if jitter_values is None:
            jitter_values = random.rand(*F.shape)
        jd = torch.from_numpy(np.ascontiguousarray(jitter_values, dtype=np.float64))[None].to(eng.device)
    auc, _, _ = eng.auc_judd(s, fx, jd, n_fix)
    return float(auc.cpu()[0])


def AUC_Borji(fixation_map, saliency_map, n_rep=100, step_size=0.1, rand_sampler=None, rand_index=None, engine=None):
    """evaluation_metrics.py:101-164.  `rand_sampler(S, F, n_rep, n_fix)` (reference hook) must return pixel
    *values* on the host, which cannot feed a device gather; pass `rand_index` (pixel indices) instead."""
    if rand_sampler is not None:
        raise ValueError("rand_sampler callbacks are host code; pass rand_index [n_fix, n_rep] pixel indices")
    eng, s, fx, F = _one('AUC_Borji', fixation_map, saliency_map, engine=engine)
    n_fix = int(F.sum())
    if n_fix == 0:
        print('no fixation to predict')
        return np.nan
    if rand_index is None:
        rand_index = random.randint(0, F.size, [n_fix, n_rep])
    rand_index = np.asarray(rand_index)
    n_rep = rand_index.shape[1]
    idx = torch.from_numpy(MetricsEngine._pad_index([rand_index], n_rep)).to(eng.device)
    mm = eng.sim_cc_nss(s, s, None)[:, 3:]          # columns 3,4 = per-frame (min, max) for normalize_range
    auc, _, _ = eng.auc_sampled(s, fx, idx, mm, 8, step_size, max(n_fix, 1))
    return float(auc.cpu()[0])


def AUC_shuffled(fixation_map, saliency_map, other_map, n_rep=100, step_size=0.1, rand_index=None, engine=None):
    """evaluation_metrics.py:167-204."""
    other_map = _dense(other_map) > 0.5
    fixation_map = _dense(fixation_map)
    if other_map.shape != fixation_map.shape:
        raise ValueError('other_map.shape != fixation_map.shape')
    F = fixation_map > 0.5
    n_fix = int(F.sum())
    if n_fix == 0:
        print('no fixation to predict')
        return np.nan
    if rand_index is None:
        rand_index = shuffled_sample_index(other_map.ravel(), n_rep, n_fix)
    return AUC_Borji(fixation_map, saliency_map, n_rep, step_size, rand_index=rand_index, engine=engine)


def _pair_stats(gtsAnn, resAnn, engine=None):
    eng = engine or default_engine()
    g = torch.from_numpy(np.ascontiguousarray(gtsAnn, dtype=np.float64))[None].to(eng.device)
    p = torch.from_numpy(np.ascontiguousarray(resAnn, dtype=np.float64))[None].to(eng.device)
    if g.shape != p.shape:
        raise ValueError("shape mismatch")
    return eng.sim_cc_nss(p, g, None).cpu().numpy()[0]


def similarity(gtsAnn, resAnn, engine=None):
    """evaluation_metrics.py:207-218."""
    return float(_pair_stats(gtsAnn, resAnn, engine)[0])


def cc(gtsAnn, resAnn, engine=None):
    """evaluation_metrics.py:221-236 (Pearson r)."""
    return float(_pair_stats(gtsAnn, resAnn, engine)[1])


def nss(fixation_map, saliency_map, engine=None):
    """Normalised scanpath saliency (not in the reference; see module docstring)."""
    eng, s, fx, F = _one('NSS', fixation_map, saliency_map, engine=engine)
    return float(eng.sim_cc_nss(s, s, fx).cpu().numpy()[0, 2])


def saliency_score_single(metric, pred_map, gt_map, fixation_map, other_map_union=None, engine=None, **rand):
    """evaluation_metrics.py:239-272."""
    if metric not in ALL_METRICS:
        raise ValueError('Unknown metric : %s' % metric)
    eng = engine or default_engine()
    kw = {}
    if 'jitter_values' in rand and rand['jitter_values'] is not None:
        kw['jitter_values'] = np.asarray(rand['jitter_values'], dtype=np.float64)[None]
    if 'rand_index' in rand and rand['rand_index'] is not None:
        kw['rand_index' if metric == 'AUC_Borji' else 'shuf_index'] = [np.asarray(rand['rand_index'])]
    fixation_map = _dense(fixation_map)
    if other_map_union is not None:
        other_map_union = _dense(other_map_union)
    if metric in ('AUC_Judd', 'AUC_Borji', 'AUC_shuffled', 'NSS') and not np.any(fixation_map > 0.5):
        if metric != 'NSS':
            print('no fixation to predict')
        return np.nan
    r = eng.evaluate_batch(_as_maps(pred_map), _as_maps(gt_map), fixation_map[None], union=other_map_union,
                           metrics=(metric,), **kw)
    return float(r[metric][0])


def other_map_union(fixation_maps, M=10):
    """evaluation_metrics.py:280-287: OR-count of M randomly chosen frames' fixations (global RNG)."""
    if len(fixation_maps) < M:
        raise AssertionError("need at least %d fixation maps" % M)
    u = np.zeros(_dense(fixation_maps[0]).shape)
    for i in random.choice(range(len(fixation_maps)), M, replace=False):
        u += (_dense(fixation_maps[i]) > 0).astype(np.int64)
    return u


def saliency_score(metric, pred_maps, gt_maps, fixation_maps, engine=None):
    """evaluation_metrics.py:275-295: NaN-propagating mean over frames, all frames in one batch on the device."""
    assert len(gt_maps) == len(pred_maps) == len(fixation_maps)
    if metric not in ALL_METRICS:
        raise ValueError('Unknown metric : %s' % metric)
    eng = engine or default_engine()
    union = other_map_union(fixation_maps)
    r = eng.evaluate_batch(np.asarray(pred_maps, dtype=np.float32), np.asarray(gt_maps, dtype=np.float32),
                           _dense(fixation_maps), union=union, metrics=(metric,))
    return float(np.mean(r[metric]))
