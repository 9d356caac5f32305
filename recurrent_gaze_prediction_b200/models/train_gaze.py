"""Training entry point — mirror of the reference's ``models/train_gaze.py`` (flags :137-155, flow :28-134).

    python -m recurrent_gaze_prediction_b200.models.train_gaze --model gaze_grcn --dataset synthetic --max_steps 20
    torchrun --nproc-per-node 8 -m recurrent_gaze_prediction_b200.models.train_gaze --model gaze_grcn ... --gpus 8

Every reference flag keeps its name and meaning.  Added: ``--dataset synthetic`` (the CRC / Hollywood2 data are not
distributable), ``--hidden --grid --n_lstm_steps --dtype --gpus --seed``.  With ``--gpus N`` (launched under
torchrun, one process per GPU) the global batch is sharded over the ranks and the flat fp32 gradient is summed with
one NCCL all-reduce before the global-norm clip and the Adam update (base.py:287-297 takes the norm of the GLOBAL
gradient)."""
from __future__ import annotations

import argparse
import logging
import os
import sys

import numpy as np

log = logging.getLogger("rgp")

MODELS = ('gaze_grcn', 'gaze_grcn77', 'gaze_lstm')


def get_model_class(name):
    """train_gaze.py:38-69 model table, restricted to the recurrent gaze-prediction path."""
    if name == 'gaze_grcn':
        from .gaze_grcn import GazePredictionGRCN as TheModel, CONSTANTS, GRUModelConfig
    elif name in ('gaze_grcn77', 'gaze_rcn77'):        # the reference imports a non-existent gaze_rcn77 (:52)
        from .gaze_grcn77 import GazePredictionGRCN as TheModel, CONSTANTS, GRUModelConfig
    elif name == 'gaze_lstm':
        from .gaze_lstm import GazePredictionLSTM as TheModel, CONSTANTS, GRUModelConfig
    elif name == 'gaze_grcn_cascade':
        from .gaze_grcn_cascade import GazePredictionGRCN as TheModel, CONSTANTS, GRUModelConfig
    else:
        raise NotImplementedError(name)
    return TheModel, CONSTANTS, GRUModelConfig


def make_session(gpus=1, device=None):
    """The reference's tf.Session becomes {'rank','world_size','allreduce'}: one process per GPU, NCCL over NVLink."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if gpus > 1 and world != gpus:
        raise RuntimeError("--gpus %d needs `torchrun --nproc-per-node %d` (WORLD_SIZE is %d)" % (gpus, gpus, world))
    sess = {"rank": rank, "world_size": world, "allreduce": None}
    dev = device or ("cuda:%d" % local)
    if dev.startswith("cuda"):
        torch.cuda.set_device(local)
        if world > 1:        # pinned staging buffers on the NUMA node of this rank's GPU (best effort)
            from .. import numa
            sess["numa"] = numa.bind_to_gpu_node(local)
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            dist.init_process_group("nccl" if dev.startswith("cuda") else "gloo")
        sess["allreduce"] = lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return sess, dev


def build_config(args, GRUModelConfig):
    """train_gaze.py:71-100."""
    config = GRUModelConfig()
    config.batch_size = 28                      # "CRC likes 28" (:75)
    config.train_dir = args.train_dir
    if args.train_tag:
        config.train_tag = args.train_tag
    config.initial_learning_rate = 0.0001
    config.max_grad_norm = 10.0
    config.use_flip_batch = True
    if args.max_grad_norm is not None:
        config.max_grad_norm = args.max_grad_norm
    if args.learning_rate is not None:
        config.initial_learning_rate = float(args.learning_rate)
    if args.learning_rate_decay is not None:
        config.learning_rate_decay = float(args.learning_rate_decay)
    if args.batch_size is not None:
        config.batch_size = int(args.batch_size)
    if args.loss_type is not None:
        config.loss_type = args.loss_type
    config.steps_per_evaluation = 100
    config.steps_per_validation = 20
    config.steps_per_checkpoint = 100
    if args.max_steps:
        config.max_steps = int(args.max_steps)
    config.rnn_state_size = args.hidden
    config.grid = args.grid
    config.n_lstm_steps = args.n_lstm_steps
    config.dtype = args.dtype
    config.seed = args.seed
    return config


def load_data_sets(args, config, CONSTANTS):
    if args.dataset != 'synthetic':
        raise NotImplementedError("dataset %r: the CRC / Hollywood2 loaders (crc_input_data_seq.py) are outside the hot "
                                  "path; use --dataset synthetic or pass any object with the next_batch() contract" % args.dataset)
    from .. import synthetic as SY
    n = max(2 * config.batch_size, 32)
    return SY.read_synthetic_data_sets(n, max(config.batch_size, 16), max(config.batch_size, 16), T=config.n_lstm_steps,
                                       grid=config.grid, GH=CONSTANTS.gazemap_height, GW=CONSTANTS.gazemap_width,
                                       dim_feature=config.dim_feature, seed=args.seed)


def train(args, session=None):
    """train_gaze.py:28-134.  `session` (optional) overrides the one derived from the torchrun environment."""
    TheModel, CONSTANTS, GRUModelConfig = get_model_class(args.model)
    if session is None:
        session, dev = make_session(args.gpus, args.device)
    else:
        dev = args.device or "cuda"
    config = build_config(args, GRUModelConfig)
    config.device = dev
    if session.get("rank", 0) == 0:
        config.dump(sys.stdout)
    log.warning('Dataset (%s) Loading ...', args.dataset)
    data_sets = load_data_sets(args, config, CONSTANTS)
    log.warning('Dataset Loading Finished ! (%d instances)', len(data_sets))
    log.warning('Start Fitting Model ...')
    model = TheModel(session, data_sets, config)
    if args.shallownet_pretrain is not None:
        raise NotImplementedError("--shallownet_pretrain: ShallowNet is not on the recurrent gaze-prediction path")
    model.fit()
    log.warning('Fitting Done. Evaluating!')
    _, scores = model.generate_and_evaluate(data_sets.test, max_instances=None)
    if session.get("rank", 0) == 0:
        for k in sorted(scores):
            print("%s : %.4f" % (k, scores[k]))
    return model, scores


def make_parser():
    parser = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    parser.add_argument('--model', required=True, type=str, help='gaze_grcn gaze_grcn77 gaze_lstm')
    parser.add_argument('--max_steps', default=None, type=int)
    parser.add_argument('--batch_norm', action='store_true')
    parser.add_argument('--batch_size', default=None, type=int)
    parser.add_argument('--gpu_fraction', default=0.48, type=float, help='accepted for compatibility; ignored')
    parser.add_argument('--train_dir', default=None, type=str)
    parser.add_argument('--train_tag', '--tag', default=None, type=str)
    parser.add_argument('--max_grad_norm', default=None, type=float)
    parser.add_argument('--learning_rate', default=None, type=float)
    parser.add_argument('--learning_rate_decay', default=None, type=float)
    parser.add_argument('--loss_type', type=str, choices=['l2', 'xentropy', 'KLD'])
    parser.add_argument('--dataset', default='crc', help='[crc, hollywood2, crcxh2, synthetic]')
    parser.add_argument('--shallownet_pretrain', default=None, type=str)
    # additions
    parser.add_argument('--hidden', default=128, type=int)
    parser.add_argument('--grid', default=7, type=int)
    parser.add_argument('--n_lstm_steps', default=42, type=int)
    parser.add_argument('--dtype', default='bf16', choices=['bf16', 'fp32'])
    parser.add_argument('--gpus', default=1, type=int)
    parser.add_argument('--seed', default=0, type=int)
    parser.add_argument('--device', default=None, type=str)
    return parser


def main(argv=None):
    logging.basicConfig(level=logging.INFO, format="%(message)s")
    args = make_parser().parse_args(argv)
    return train(args)


if __name__ == '__main__':
    main()
