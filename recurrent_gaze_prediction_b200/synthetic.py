"""Synthetic clips with the reference loader's tuple contract.

Replaces ``crc_input_data_seq.read_crc_data_sets`` (reference
``crc_input_data_seq.py:123-156``: ``next_batch(B) -> (images, gazemaps,
fixationmaps, c3ds, pupils, clipnames)``) for runs without the CRC/Hollywood2
data, which are not distributable.  Generators follow SURVEY.md 8(d):
post-ReLU-like C3D conv5b features, Gaussian-blob gaze maps (sigma 2 px at
49x49, 0.3 at 7x7 — ``crc_input_data_seq.py:225-233``), binary fixation maps
drawn from the gaze density.  Host-side NumPy only; no device work here.
"""
from __future__ import annotations

import numpy as np


def make_c3d(rng, B, T, dim_feature=1024, grid=7, dtype=np.float32):
    """c3d[B,T,1024,G,G] = max(0, N(0,1)) (conv5b is post-ReLU; ~50% zeros)."""
    x = rng.standard_normal(size=(B, T, dim_feature, grid, grid)).astype(dtype)
    np.maximum(x, 0, out=x)
    return x


def make_embedded(rng, B, T, grid=7, dim_proj=512, dtype=np.float32):
    """'L1' features entering after the 1024->512 projection: emb[B,T,G,G,512] ~ N(0,1.8^2)."""
    return (1.8 * rng.standard_normal(size=(B, T, grid, grid, dim_proj))).astype(dtype)


def make_gazemaps(rng, B, T, GH=49, GW=49, dtype=np.float32, normalize=False):
    """1-3 isotropic Gaussians per frame, min-max normalised (loader), optionally sum=1."""
    sigma = 2.0 if GH >= 49 else (0.3 if GH <= 7 else 0.6)
    yy, xx = np.mgrid[0:GH, 0:GW].astype(np.float64)
    out = np.zeros((B, T, GH, GW), dtype=np.float64)
    centres = []
    for b in range(B):
        for t in range(T):
            k = rng.randint(1, 4)
            cs = np.stack([rng.uniform(0, GH - 1, size=k), rng.uniform(0, GW - 1, size=k)], axis=1)
            m = np.zeros((GH, GW))
            for cy, cx in cs:
                m += np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * sigma * sigma))
            m = (m - m.min()) / (m.max() - m.min())
            out[b, t] = m
            centres.append(cs)
    if normalize:
        out /= out.sum(axis=(2, 3), keepdims=True)
    return out.astype(dtype), centres


def make_fixationmaps(rng, gazemaps, H=49, W=49, max_fix=20):
    """Binary uint8 [B,T,H,W] with 1..max_fix fixations per frame drawn from the gaze density
    (nearest-neighbour up-scaling of the density when H,W differ from the gaze-map size)."""
    B, T, GH, GW = gazemaps.shape
    out = np.zeros((B, T, H, W), dtype=np.uint8)
    ry = (np.arange(H) * GH // H).astype(np.int64)
    rx = (np.arange(W) * GW // W).astype(np.int64)
    for b in range(B):
        for t in range(T):
            dens = gazemaps[b, t].astype(np.float64)[ry][:, rx].ravel()
            dens = dens / dens.sum()
            n = rng.randint(1, max_fix + 1)
            idx = rng.choice(H * W, size=n, replace=True, p=dens)
            out[b, t].reshape(-1)[idx] = 1
    return out


class SyntheticDataSet(object):
    """Duck-type of CRCDataSet (crc_input_data_seq.py:84-156): __len__, next_batch wrap-around."""

    def __init__(self, n_clips, T=35, seed=0, grid=7, GH=49, GW=49, image_hw=98,
                 fix_hw=None, with_images=False, dim_feature=1024):
        rng = np.random.RandomState(seed)
        self.c3ds = make_c3d(rng, n_clips, T, dim_feature, grid).reshape(n_clips, T, -1)
        self.gazemaps, _ = make_gazemaps(rng, n_clips, T, GH, GW)
        fh, fw = fix_hw or (GH, GW)
        self.fixationmaps = make_fixationmaps(rng, self.gazemaps, fh, fw)
        ih = image_hw if with_images else 1
        self.images = np.zeros((n_clips, T, ih, ih, 3), dtype=np.float32)
        self.pupils = np.zeros((n_clips, T), dtype=np.float32)
        self.clipnames = ["synthetic_%05d" % i for i in range(n_clips)]
        self.epochs_completed = 0
        self.index_in_epoch = 0

    def __len__(self):
        return len(self.c3ds)

    def image_count(self):
        return len(self.c3ds)

    def next_batch(self, batch_size):
        start = self.index_in_epoch
        self.index_in_epoch += batch_size
        if self.index_in_epoch > self.image_count():
            self.epochs_completed += 1
            start = 0
            self.index_in_epoch = batch_size
            assert batch_size <= self.image_count()
        end = self.index_in_epoch
        s = slice(start, end)
        return (self.images[s], self.gazemaps[s].copy(), self.fixationmaps[s].copy(),
                self.c3ds[s].copy(), self.pupils[s], self.clipnames[s])


class SyntheticDataSplits(object):
    """Duck-type of CRCDataSplits: .train / .valid / .test and len()."""

    def __init__(self, n_train=64, n_valid=16, n_test=16, **kw):
        seed = kw.pop("seed", 0)
        self.train = SyntheticDataSet(n_train, seed=seed, **kw)
        self.valid = SyntheticDataSet(n_valid, seed=seed + 1, **kw)
        self.test = SyntheticDataSet(n_test, seed=seed + 2, **kw)

    def __len__(self):
        return len(self.train) + len(self.valid) + len(self.test)


def read_synthetic_data_sets(n_train=64, n_valid=16, n_test=16, **kw):
    return SyntheticDataSplits(n_train, n_valid, n_test, **kw)
