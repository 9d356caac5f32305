"""Evaluation entry point — mirror of the reference's ``models/evaluate_gaze.py`` (flags :289-303; flow
``load_model`` :48-107, ``handle_frame`` :117-157, ``run_evaluation`` :171-226, ``main`` :243-286).

The reference scores frame by frame in a joblib fork pool (``-j``); here all sampled frames are scored in batched
launches of the float64 metric kernels (``-j`` is accepted and ignored).  The per-frame random draws (the 10-frame
fixation union, Borji negatives, Judd jitter, shuffled permutations) come from the global ``numpy.random`` state in
``handle_frame``'s order.  Image dumps (``scipy.misc.imsave``) are left out; ``overall.txt`` keeps its format."""
from __future__ import annotations

import argparse
import logging
import os
from collections import OrderedDict

import numpy as np

from .. import evaluation_metrics as EM
from .train_gaze import get_model_class, make_session

log = logging.getLogger("rgp")
HANDLE_FRAME_METRICS = ('sim', 'cc', 'AUC_Borji', 'AUC_Judd', 'AUC_shuffled')      # evaluate_gaze.py:134


def load_model(args, session=None, device=None):
    """evaluate_gaze.py:48-107."""
    TheModel, CONSTANTS, GRUModelConfig = get_model_class(args.model)
    config = GRUModelConfig()
    config.batch_size = args.batch_size or 14
    config.loss_type = args.loss_type if args.loss_type is not None else 'xentropy'
    config.rnn_state_size, config.grid = args.hidden, args.grid
    config.n_lstm_steps, config.dtype = args.n_lstm_steps, args.dtype
    config.trainable = False
    if device is not None:
        config.device = device
    from .. import synthetic as SY
    empty = SY.SyntheticDataSplits.__new__(SY.SyntheticDataSplits)
    empty.train = empty.valid = empty.test = []
    return TheModel(session, empty, config)


def reload_checkpoint(model, checkpoint_path):
    """evaluate_gaze.py:160-168."""
    model.load_model_from_checkpoint_file(os.path.abspath(checkpoint_path))
    log.warning('Model Loading Done!!')


def frame_unions(fixationmaps, M=10):
    """handle_frame's per-frame other_map_union (evaluate_gaze.py:124-131): M random OTHER-frame picks per frame."""
    out = []
    for fm in fixationmaps:
        u = np.zeros(np.asarray(fm).shape, dtype=np.uint8)
        for oth in np.random.choice(range(len(fixationmaps)), M, replace=False):
            other = np.asarray(fixationmaps[oth])
            if other.shape != u.shape:
                other = EM.resize_onehot_tensor_sparse(other, u.shape)
            u += (other > 0).astype(np.uint8)
        out.append(u)
    return out


def score_frames(pred_gazemaps, gt_gazemaps, fixationmaps, engine=None, rng="numpy"):
    """All handle_frame metrics for all frames.  Frames are grouped by fixation-map shape (original-scale fixation
    maps may differ per clip) and by union; sim / cc / Borji / Judd go through in one batch per shape."""
    eng = engine or EM.default_engine()
    n = len(pred_gazemaps)
    unions = frame_unions(fixationmaps)
    scores = OrderedDict((m, np.full(n, np.nan)) for m in HANDLE_FRAME_METRICS)
    shapes = OrderedDict()
    for i, fm in enumerate(fixationmaps):
        shapes.setdefault(np.asarray(fm).shape, []).append(i)
    for shp, idx in shapes.items():
        pred = np.asarray([pred_gazemaps[i] for i in idx], dtype=np.float32)
        gt = np.asarray([gt_gazemaps[i] for i in idx], dtype=np.float32)
        fix = np.asarray([fixationmaps[i] for i in idx])
        r = eng.evaluate_batch(pred, gt, fix, metrics=('sim', 'cc', 'AUC_Borji', 'AUC_Judd'), rng=rng)
        for m in r:
            scores[m][idx] = r[m]
        for j, i in enumerate(idx):           # the shuffled AUC has a different union per frame
            if (fix[j] > 0.5).any():
                r = eng.evaluate_batch(pred[j:j + 1], gt[j:j + 1], fix[j:j + 1], union=unions[i],
                                       metrics=('AUC_shuffled',), rng=rng, seed=i)
                scores['AUC_shuffled'][i] = r['AUC_shuffled'][0]
    return scores


def run_evaluation(model, args, data_sets, out_dir):
    """evaluate_gaze.py:171-226."""
    assert out_dir is not None
    T = model.n_lstm_steps
    ret = model.generate(data_sets.valid, max_instances=int(np.divide(args.num_frames, T, dtype=float) + 1))
    pred, gt, fix = np.asarray(ret['pred_gazemap_list']), np.asarray(ret['gt_gazemap_list']), ret['fixationmap_list']
    n_images = len(pred)
    assert n_images == len(gt) == len(fix)
    scores = score_frames(pred, gt, fix, engine=EM.MetricsEngine(device=model.engine.device, lib=model._lib),
                          rng="device" if getattr(args, "fast_rng", False) else "numpy")
    os.makedirs(out_dir, exist_ok=True)
    overall_txt_path = os.path.join(out_dir, 'overall.txt')
    with open(overall_txt_path, 'w') as fp:
        for metric, score_list in scores.items():
            mean_score = np.mean(score_list)
            log.info("Average %s : %.4f" % (metric, mean_score))
            fp.write("Average %s : %.4f\n" % (metric, mean_score))
            fp.write(' '.join('%.3f' % s for s in score_list) + ' \n')
    log.warning("Dumped at %s", overall_txt_path)
    return scores


def get_out_dir(dataset, checkpoint_path):
    """evaluate_gaze.py:229-241."""
    try:
        checkpoint_step = os.path.splitext(os.path.basename(checkpoint_path))[0].split('-')[-1]
    except Exception:
        checkpoint_step = 'latest'
    out_dir = os.path.abspath(os.path.join(checkpoint_path, '../../generated-{}-{}'.format(dataset, checkpoint_step)))
    os.makedirs(out_dir, exist_ok=True)
    return out_dir


def main(argv=None):
    logging.basicConfig(level=logging.INFO, format="%(message)s")
    args = make_parser().parse_args(argv)
    if not args.embed and not args.checkpoint_path:
        raise ValueError('checkpoint_path needed')
    if args.embed:
        raise NotImplementedError("--embed (an IPython shell) is not provided")
    if args.dataset != 'synthetic':
        raise NotImplementedError("dataset %r: only --dataset synthetic ships (see train_gaze.py)" % args.dataset)
    session, dev = make_session(1, args.device)
    model = load_model(args, session, dev)
    from .. import synthetic as SY
    n_clips = int(args.num_frames / model.n_lstm_steps) + 1 + model.batch_size
    fix_hw = (360, 480) if args.fixation_original_scale else None
    data_sets = SY.read_synthetic_data_sets(model.batch_size, n_clips, model.batch_size, T=model.n_lstm_steps,
                                            grid=args.grid, GH=model.gazemap_height, GW=model.gazemap_width,
                                            fix_hw=fix_hw, seed=args.seed)
    reload_checkpoint(model, args.checkpoint_path)
    out_dir = get_out_dir(args.dataset, args.checkpoint_path)
    return run_evaluation(model, args, data_sets, out_dir)


def make_parser():
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    parser.add_argument('--model', required=True, type=str)
    parser.add_argument('--batch_size', default=None, type=int)
    parser.add_argument('--dataset', default='crc', choices=['hollywood2', 'crc', 'synthetic'])
    parser.add_argument('--debug', action='store_true')
    parser.add_argument('--gpu_fraction', default=0.48, type=float, help='accepted for compatibility; ignored')
    parser.add_argument('--checkpoint_path', required=False, type=str)
    parser.add_argument('--loss_type', type=str, choices=['l2', 'xentropy'])
    parser.add_argument('--num_frames', type=int, default=1500)
    parser.add_argument('-j', '--jobs', type=int, default=10, help='accepted for compatibility; frames are batched on the GPU')
    parser.add_argument('--embed', action='store_true')
    parser.add_argument('--fast_rng', action='store_true',
                        help="draw the AUC jitter / negatives on the device (statistically equivalent, not NumPy's global stream)")
    # additions
    parser.add_argument('--hidden', default=128, type=int)
    parser.add_argument('--grid', default=7, type=int)
    parser.add_argument('--n_lstm_steps', default=42, type=int)
    parser.add_argument('--dtype', default='bf16', choices=['bf16', 'fp32'])
    parser.add_argument('--seed', default=0, type=int)
    parser.add_argument('--device', default=None, type=str)
    parser.add_argument('--fixation_original_scale', action='store_true',
                        help='score against 360x480 fixation maps (evaluate_gaze.py:263), bicubic-resizing the 49x49 maps')
    return parser


if __name__ == '__main__':
    main()
