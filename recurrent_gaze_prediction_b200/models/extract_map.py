"""Gaze-map extraction for downstream models — mirror of the reference's ``models/extract_map.py``
(``:52-82`` model set-up with ``n_lstm_steps = 105``, ``:114-205`` ``generate_batches``: zero-pad short clips, cut long
ones, ``:210-229`` inference and the two ``.npy`` dumps per clip: ``<name>.gazemap.49.npy`` [len,49,49] and the 7x7
average-pooled ``<name>.gazemap.npy`` [len,7,7]).

The reference feeds ``model.c3d_input`` / ``model.frame_images`` to ``session.run`` directly and notes
"TODO INSTEAD, MAKE INFERENCE() function in the model" (``:209``); here that call is ``model.predict``.  Long clips run
through the engine ``--chunk_steps`` frames at a time with the recurrent state carried from chunk to chunk, so the
activation buffers are sized for one chunk while the clip may be arbitrarily long.
Frame images are not read: the GRCN graph never consumes them (``gaze_grcn.py:173-376``).

C3D features come from ``--c3d_dir`` (one ``<clip>.npy`` [len,1024,7,7] per clip; the reference's hdf5 tree of LSMDC
needs h5py, which this image does not have) or from ``--dataset synthetic``.
"""
from __future__ import annotations

import argparse
import logging
import os

import numpy as np

from .train_gaze import get_model_class, make_session

log = logging.getLogger("rgp")


def avg_pool(gazemaps, k=7):
    """extract_map.py:34-44 `avg_pool` (49x49 -> 7x7 block means), as the host post-processing it is there."""
    g = np.asarray(gazemaps)
    n, H, W = g.shape
    assert H % k == 0 and W % k == 0
    return g.reshape(n, H // k, k, W // k, k).mean(axis=(2, 4))


def fit_length(c3d, T):
    """extract_map.py:139-176: zero-pad a short clip to T frames, cut a long one; returns (c3d[T,...], kept length)."""
    c3d = np.asarray(c3d, dtype=np.float32)
    n = c3d.shape[0]
    if n < T:
        c3d = np.pad(c3d, [(0, T - n)] + [(0, 0)] * (c3d.ndim - 1), 'constant', constant_values=np.float32(0.0))
    elif n > T:
        log.warning('Too long. c3d_len = %d, rnn steps = %d', n, T)
        c3d, n = c3d[:T], T
    return c3d, n


def generate_batches(clips, B, T, feat_shape=(1024, 7, 7)):
    """extract_map.py:114-205: yields (names, batch_c3d [B,T,...], lengths); the last batch is padded with empty slots
    (name None, length 0)."""
    batch = np.zeros((B, T) + tuple(feat_shape), dtype=np.float32)
    names, lengths, b = [None] * B, [0] * B, 0
    for name, c3d in clips:
        batch[b], lengths[b] = fit_length(c3d, T)
        names[b] = name
        b += 1
        if b >= B:
            yield list(names), batch, list(lengths)
            batch = np.zeros_like(batch)
            names, lengths, b = [None] * B, [0] * B, 0
    if b > 0:
        yield list(names), batch, list(lengths)


def extract(model, clips, out_dir=None):
    """Runs every clip of `clips` (iterable of (name, c3d [len,1024,7,7])) through `model.predict`; returns
    {name: (gaze49 [len,49,49], gaze7 [len,7,7])} and, with `out_dir`, writes the reference's two files per clip."""
    if model.world_size != 1:
        raise NotImplementedError("extract() drives one GPU; shard the clip list across processes instead")
    B, T = model.batch_size, model.n_lstm_steps
    G = model.engine.G
    out = {}
    for names, batch_c3d, lengths in generate_batches(clips, B, T, (model.dim_feature, G, G)):
        gazes = model.predict(batch_c3d)
        for name, length, gaze in zip(names, lengths, gazes):
            if name is None:
                continue
            g49 = np.array(gaze[:length])
            g7 = avg_pool(g49) if g49.shape[1] % 7 == 0 and g49.shape[1] > 7 else g49
            out[name] = (g49, g7)
            if out_dir is not None:
                d = os.path.join(out_dir, name)
                os.makedirs(d, exist_ok=True)
                np.save(os.path.join(d, '%s.gazemap.49.npy' % name), g49)
                np.save(os.path.join(d, '%s.gazemap.npy' % name), g7)
                log.info('%s : saved length = %d', name, length)
    return out


def load_model(args, session=None, device=None):
    """extract_map.py:52-82."""
    TheModel, CONSTANTS, GRUModelConfig = get_model_class(args.model)
    config = GRUModelConfig()
    config.batch_size = args.batch_size or 14
    config.train_dir = None
    config.n_lstm_steps = args.n_lstm_steps
    config.chunk_steps = args.chunk_steps
    config.loss_type = args.loss_type if args.loss_type is not None else 'xentropy'
    config.rnn_state_size, config.grid, config.dtype = args.hidden, args.grid, args.dtype
    config.trainable = False
    if device is not None:
        config.device = device
    from .. import synthetic as SY
    empty = SY.SyntheticDataSplits.__new__(SY.SyntheticDataSplits)
    empty.train = empty.valid = empty.test = []
    model = TheModel(session, empty, config)
    if args.checkpoint_path:
        model.load_model_from_checkpoint_file(os.path.abspath(args.checkpoint_path))
    return model


def iter_clips(args):
    if args.dataset == 'synthetic':
        from .. import synthetic as SY
        rng = np.random.RandomState(args.seed)
        for i in range(args.num_clips):
            n = int(rng.randint(max(1, args.n_lstm_steps // 3), args.n_lstm_steps + 10))
            yield 'clip%04d' % i, SY.make_c3d(rng, 1, n, grid=args.grid)[0]
    else:
        if not args.c3d_dir:
            raise ValueError("--c3d_dir needed (one <clip>.npy [len,1024,7,7] per clip)")
        for fn in sorted(os.listdir(args.c3d_dir)):
            if fn.endswith('.npy'):
                name = fn[:-4]
                if os.path.exists(os.path.join(args.out_dir, name)):
                    log.warning('Skipped - already exists %s', name)     # extract_map.py:128-131
                    continue
                yield name, np.load(os.path.join(args.c3d_dir, fn))


def make_parser():
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    parser.add_argument('--model', required=True, type=str)
    parser.add_argument('--batch_size', default=None, type=int)
    parser.add_argument('--checkpoint_path', required=False, type=str)
    parser.add_argument('--loss_type', type=str, choices=['l2', 'xentropy'])
    parser.add_argument('--gpu_fraction', default=0.48, type=float, help='accepted for compatibility; ignored')
    # additions
    parser.add_argument('--dataset', default='lsmdc', choices=['lsmdc', 'synthetic'])
    parser.add_argument('--c3d_dir', default=None, type=str)
    parser.add_argument('--out_dir', default='gazemap_cowork', type=str)
    parser.add_argument('--num_clips', default=28, type=int)
    parser.add_argument('--n_lstm_steps', default=105, type=int)
    parser.add_argument('--chunk_steps', default=35, type=int)
    parser.add_argument('--hidden', default=128, type=int)
    parser.add_argument('--grid', default=7, type=int)
    parser.add_argument('--dtype', default='bf16', choices=['bf16', 'fp32'])
    parser.add_argument('--seed', default=0, type=int)
    parser.add_argument('--device', default=None, type=str)
    return parser


def main(argv=None, session=None):
    logging.basicConfig(level=logging.INFO, format="%(message)s")
    args = make_parser().parse_args(argv)
    if session is None:
        session, dev = make_session(1, args.device)
    else:
        dev = args.device
    model = load_model(args, session, dev)
    log.info('Batch Size : %d', model.batch_size)
    return extract(model, iter_clips(args), args.out_dir)


if __name__ == '__main__':
    main()
