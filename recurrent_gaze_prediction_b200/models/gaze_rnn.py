"""Recurrent gaze-prediction model base — mirror of the reference's ``models/gaze_rnn.py``.

``GazePredictionGRU`` keeps the reference's constructor / attribute / method contract (``gaze_rnn.py:66-105`` ctor,
``:108-209`` build_model, ``:363-408`` create_loss_and_summary, ``:435-478`` learning rate + train op, ``:483-565``
single_step, ``:568-650`` generate, ``:653-679`` evaluate / generate_and_evaluate).  The TF session is replaced by
a :class:`GazeEngine`: "placeholders" (``c3d_input``, ``frame_images``, ``gt_gazemap``) are device buffers that
``single_step`` / ``generate`` fill from pinned host memory, and ``session.run`` is one forward (+ backward +
clip/Adam) pass of hand-written kernels on the model's CUDA stream.

The class named GazePredictionGRU in the reference is itself the *dense* GRU variant (parity
unpinned at the tf.nn.rnn_cell.GRUCell boundary): instantiated directly it builds that network on
:class:`DenseGruEngine` (forward / loss / generate / evaluate; no training).  The GRCN / GRCN77 / LSTM subclasses
carry the hot path.
"""
from __future__ import annotations

import logging
import time

import numpy as np
import torch

from ..engine import GazeEngine, LOSS_TYPES, StepGraph
from .. import lib as L
from .. import evaluation_metrics as EM
from .base import BaseModelConfig, ModelBase
from .model_util import init_parameters, normalize_probability_map

log = logging.getLogger("rgp")


class _E(dict):
    """attribute dict (the reference uses easydict)"""
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


CONSTANTS = _E(image_width=98, image_height=98, gazemap_width=49, gazemap_height=49,
               saliencymap_width=49, saliencymap_height=49)        # gaze_rnn.py:34-40


class GRUModelConfig(BaseModelConfig):
    """gaze_rnn.py:44-61, plus the knobs the reference hard-codes inside its graph builders (hidden width 128
    ``gaze_grcn.py:211``, projection 512 ``:204``, grid 7) and the B200 execution options."""

    def __init__(self):
        super(GRUModelConfig, self).__init__()
        self.n_lstm_steps = 42
        self.batch_size = 7
        self.dim_feature = 1024
        self.dim_sal = 1024 * 49
        self.dim_sal_proj = 1024
        self.optimization_method = 'adam'
        self.loss_type = 'xentropy'
        self.use_flip_batch = True
        # --- extensions (defaults reproduce the reference) ---
        self.rnn_state_size = 128
        self.dim_cnn_proj = 512
        self.grid = 7
        self.dtype = 'bf16'           # 'bf16': tcgen05 engine, fp32 accumulate; 'fp32': CUDA-core engine
        self.device = 'cuda'
        self.seed = 0
        self.learning_rate_decay_steps = 500      # gaze_rnn.py:440 hard-codes 500
        # execution options of the B200 path
        self.cuda_graph = True        # one CUDA-graph submission per step (fwd [+ BPTT + clip/Adam]) instead of ~80-900 launches
        self.prefetch = True          # fit(): batch k+1 is prepared on the host and uploaded while step k runs
        self.feature_dtype = 'fp32'   # dtype of the fed C3D features: 'fp32' (reference) or 'bf16' (half the H2D bytes;
                                      # bit-identical results on the bf16 engine, which rounds the features anyway)


def _shard(x, rank, world):
    n = len(x) // world
    return x[rank * n:(rank + 1) * n]


class GazePredictionGRU(ModelBase):
    model_name = None            # engine model key, set by subclasses
    train_mode_default = True

    def __init__(self, session, data_sets, config=None, gazemap_height=None, gazemap_width=None):
        config = config if config is not None else GRUModelConfig()
        self.session = session            # opaque: None, or a dict {'rank':, 'world_size':, 'allreduce':, 'lib':}
        self.data_sets = data_sets
        self.config = config
        super(GazePredictionGRU, self).__init__(config)
        self.batch_size = config.batch_size
        self.n_lstm_steps = config.n_lstm_steps
        self.dim_feature = config.dim_feature
        self.dim_sal = config.dim_sal
        self.dim_sal_proj = config.dim_sal_proj
        self.dim_cnn_proj = getattr(config, "dim_cnn_proj", 512)
        self.gazemap_height = gazemap_height or self._default_gazemap_hw()[0]
        self.gazemap_width = gazemap_width or self._default_gazemap_hw()[1]
        self.image_height = CONSTANTS.image_height
        self.image_width = CONSTANTS.image_width
        sess = session if isinstance(session, dict) else {}
        self.rank = int(sess.get("rank", 0))
        self.world_size = int(sess.get("world_size", 1))
        self._allreduce = sess.get("allreduce")
        self._lib = sess.get("lib")
        if self.batch_size % self.world_size:
            raise ValueError("batch_size %d is not divisible by the %d data-parallel ranks" % (self.batch_size, self.world_size))
        self.local_batch = self.batch_size // self.world_size
        self.build_model()
        self.build_train_op()

    def _default_gazemap_hw(self):
        return CONSTANTS.gazemap_height, CONSTANTS.gazemap_width

    # ---------------------------------------------------------------- graph = engine + device "placeholders"
    def build_model(self):
        """gaze_rnn.py:108-209: inputs, network, loss."""
        if self.model_name is None:
            return self._build_dense_model()
        c = self.config
        B, T, G = self.local_batch, self.n_lstm_steps, getattr(c, "grid", 7)
        dev = torch.device(getattr(c, "device", "cuda"))
        dtype = 1 if getattr(c, "dtype", "bf16") == "bf16" else 0
        train = bool(getattr(c, "trainable", True))
        # chunk_steps: long clips (extract_map.py:65 runs T = 105) go through the engine `chunk_steps` frames at a time
        # with the recurrent state carried between chunks; activations are sized for one chunk.  Inference only.
        self.chunk_steps = getattr(c, "chunk_steps", None) or None
        Tc = self.chunk_steps or T
        if self.chunk_steps is not None:
            if train:
                raise ValueError("chunk_steps is an inference option (set config.trainable = False)")
            if T % Tc != 0:
                raise ValueError("n_lstm_steps=%d is not a multiple of chunk_steps=%d" % (T, Tc))
        self.engine = GazeEngine(self.model_name, B=B, T=Tc, G=G, Ch=getattr(c, "rnn_state_size", 128),
                                 Cin=self.dim_cnn_proj, Cfeat=c.dim_feature, dtype=dtype, device=dev, lib=self._lib,
                                 train=train, T_params=T)
        GH = self.engine.GH
        if (self.gazemap_height, self.gazemap_width) != (GH, GH):
            raise ValueError("%s predicts %dx%d gaze maps, not %dx%d" % (self.model_name, GH, GH, self.gazemap_height,
                                                                         self.gazemap_width))
        self.engine.load_tf_params(init_parameters(self.model_name, hidden=self.engine.Ch, T=T, grid=G,
                                                   dim_proj=self.dim_cnn_proj, dim_feature=c.dim_feature,
                                                   seed=getattr(c, "seed", 0)))
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        fdt = {'fp32': torch.float32, 'bf16': torch.bfloat16}[getattr(c, "feature_dtype", 'fp32')]
        self.frame_images = None                                     # unused by the RGP path (gaze_grcn.py:173-376 never reads it)
        self.c3d_input = torch.zeros(B, T, c.dim_feature, G, G, dtype=fdt, device=dev)   # gaze_rnn.py:118-121
        self.gt_gazemap = z(B, T, GH, GH)                            # gaze_rnn.py:122-126
        self.dropout_keep_prob = 1.0                                 # never fires in GRCN / LSTM
        if self.chunk_steps is None:
            self.predicted_gazemaps_logit = self.engine.logits       # gaze_rnn.py:156
            probs = self.engine.probs
        else:
            self.predicted_gazemaps_logit, probs = z(B, T, GH, GH), z(B, T, GH, GH)
        self._probs = probs
        self.predicted_gazemaps = probs if c.loss_type in ('xentropy', 'KLD') else self.predicted_gazemaps_logit
        self.loss = self.engine.loss
        if c.loss_type not in LOSS_TYPES:
            raise NotImplementedError(str(c.loss_type))             # gaze_rnn.py:405-406
        self._alloc_staging(dev, fdt)

    def _alloc_staging(self, dev, fdt=torch.float32):
        """pinned host staging for batches that arrive as ordinary (pageable) arrays, the pinned loss slot, and the
        double-buffered upload path used by fit() (pipeline.PrefetchFeeder)"""
        pin = dev.type == "cuda"
        self._c3d_host = torch.zeros(tuple(self.c3d_input.shape), dtype=fdt, pin_memory=pin)
        self._gt_host = torch.zeros(tuple(self.gt_gazemap.shape), dtype=torch.float32, pin_memory=pin)
        self._loss_host = torch.zeros(1, dtype=torch.float32, pin_memory=pin)
        self._loss_event = torch.cuda.Event() if pin else None
        self._loss_pending = False
        self._last_loss = float("nan")
        self._graphs = {}
        self._feeder = None
        self._gather_buf = None

    def _build_dense_model(self):
        """The class as the reference defines it: the dense GRU variant (gaze_rnn.py:211-361) —
        forward, loss, BPTT + clip/Adam (eager launches), generate / evaluate.  Dropout on the embedded features is live
        while training (keep_prob 0.5, gaze_rnn.py:302-303,529) and off otherwise."""
        from ..dense_engine import DenseGruEngine
        from .model_util import init_parameters_dense
        c = self.config
        B, T, G = self.local_batch, self.n_lstm_steps, getattr(c, "grid", 7)
        dev = torch.device(getattr(c, "device", "cuda"))
        dtype = 1 if getattr(c, "dtype", "bf16") == "bf16" else 0
        GH, GW = self.gazemap_height, self.gazemap_width
        if GH != GW:
            raise ValueError("square gaze maps only")
        train = bool(getattr(c, "trainable", True))
        self.dim_cnn_proj = getattr(c, "dim_cnn_proj_dense", 32)              # gaze_rnn.py:239 (XXX FIXME there)
        self.chunk_steps = None
        self.engine = DenseGruEngine(B=B, T=T, G=G, Cp=self.dim_cnn_proj, Cfeat=c.dim_feature, GH=GH, dtype=dtype,
                                     device=dev, lib=self._lib, train=train)
        self.engine.load_tf_params(init_parameters_dense(dim_proj=self.dim_cnn_proj, grid=G, dim_feature=c.dim_feature,
                                                         GH=GH, GW=GW, seed=getattr(c, "seed", 0)))
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        self.frame_images = None
        self.c3d_input = z(B, T, c.dim_feature, G, G)
        self.gt_gazemap = z(B, T, GH, GW)
        self.dropout_keep_prob = 1.0       # gaze_rnn.py:128 placeholder_with_default(1.0); 0.5 is fed only while training (:529)
        self.predicted_gazemaps_logit = self.engine.logits
        self._probs = self.engine.probs
        self.predicted_gazemaps = self._probs if c.loss_type in ('xentropy', 'KLD') else self.engine.logits
        self.loss = self.engine.loss
        if c.loss_type not in LOSS_TYPES:
            raise NotImplementedError(str(c.loss_type))
        self._alloc_staging(dev)

    @staticmethod
    def create_loss_and_summary(predicted_gazemaps, gt_gazemap, loss_type='xentropy', lib=None):
        """gaze_rnn.py:363-408: scalar loss of logits [B,T,GH,GW] against gt (device tensors).  'KLD' (broken in
        the reference, :395-403) is xent - H(gt), which has the same gradient."""
        if loss_type not in LOSS_TYPES:
            raise NotImplementedError(str(loss_type))
        lib = lib or L.get_lib()
        B, T, GH, GW = predicted_gazemaps.shape
        assert tuple(gt_gazemap.shape) == (B, T, GH, GW)
        dev = predicted_gazemaps.device
        fl = torch.zeros(B * T, dtype=torch.float32, device=dev)
        out = torch.zeros(1, dtype=torch.float32, device=dev)
        s = torch.cuda.current_stream(dev).cuda_stream if dev.type == "cuda" else 0
        lib.rgp_softmax_xent_fwd_bwd(predicted_gazemaps.contiguous().data_ptr(), gt_gazemap.contiguous().data_ptr(), 0, 0,
                                     fl.data_ptr(), B * T, GH * GW, 1.0 / (B * T), LOSS_TYPES[loss_type], s)
        lib.rgp_reduce_sum(fl.data_ptr(), B * T, out.data_ptr(), s)
        return out

    # ---------------------------------------------------------------- optimiser (gaze_rnn.py:435-478, base.py:262-297)
    def _learning_rate_at(self, step):
        """tf.train.exponential_decay(lr0, global_step, 500, decay, staircase=True) (gaze_rnn.py:439-443)."""
        c = self.config
        return self.initial_learning_rate * self._lr_scale * \
            self.learning_rate_decay ** (step // getattr(c, "learning_rate_decay_steps", 500))

    def build_train_op(self):
        method = getattr(self.config, "optimization_method", 'adam')
        if method not in ('adam', 'rmsprop', 'sgd'):
            raise ValueError('Invalid optimization method!')                  # base.py:268-275
        if getattr(self.engine, "train", False):
            # base.py:268-273: Adam / RMSProp(momentum 0.9) / Momentum(0.9), all on the globally clipped gradient
            self.engine.set_optimizer(method)
        self.global_step = 0
        self.learning_rate = self.current_learning_rate
        self.train_op = self._train_step if getattr(self.engine, "train", False) else None
        self.no_op = lambda: None        # the ShallowNet train-op of the reference is a no_op (gaze_rnn.py:459-466)

    def _opt_kwargs(self):
        return dict(lr0=self.initial_learning_rate * self._lr_scale, decay=self.learning_rate_decay,
                    clip=self.max_grad_norm, decay_steps=int(getattr(self.config, "learning_rate_decay_steps", 500)))

    def _train_step(self):
        self.engine.backward()
        self.engine.optimizer_step(allreduce=self._allreduce, **self._opt_kwargs())
        self.global_step = self.engine.step

    # ---------------------------------------------------------------- feeding ("session.run")
    @staticmethod
    def _as_host_tensor(x, dtype=None):
        """NumPy array or CPU tensor -> contiguous CPU tensor (no copy when it already is one of the right dtype)"""
        t = x if isinstance(x, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(x))
        if dtype is not None and t.dtype != dtype:
            t = t.to(dtype)
        return t.contiguous()

    def _stage(self, batch_c3d, batch_maps, stage=None, local=False):
        """host batch (NumPy arrays or CPU tensors) -> this rank's shard as PINNED host tensors of the placeholders'
        dtypes.  The batch is the GLOBAL one (batch_size clips, as the reference's next_batch returns it) of which this
        rank keeps its slice, or — local=True — already this rank's own `local_batch` clips (a data-parallel loader that
        reads only its shard).  A shard that already lives in pinned memory with the right dtype is used in place (no
        staging copy: the H2D engine reads it directly); anything else goes through pinned staging buffers."""
        G = self.engine.G
        nb = self.local_batch if local else self.batch_size
        c3d = batch_c3d if isinstance(batch_c3d, torch.Tensor) else np.asarray(batch_c3d)
        c3d = c3d.reshape(nb, -1, self.dim_feature, G, G)
        maps = batch_maps if isinstance(batch_maps, torch.Tensor) else np.asarray(batch_maps, dtype=np.float32)
        if len(maps) != nb:
            raise ValueError("gaze maps for %d clips, features for %d" % (len(maps), nb))
        if not local:
            c3d = _shard(c3d, self.rank, self.world_size)
            maps = _shard(maps, self.rank, self.world_size)
        if c3d.shape[1] != self.n_lstm_steps:
            raise ValueError("clip length %d != n_lstm_steps %d" % (c3d.shape[1], self.n_lstm_steps))
        out = []
        for x, buf in zip((c3d, maps), stage or (self._c3d_host, self._gt_host)):
            t = x if isinstance(x, torch.Tensor) else torch.from_numpy(x)
            if t.dtype == buf.dtype and t.is_contiguous() and buf.is_pinned() and t.is_pinned():
                out.append(t.view(buf.shape))
            else:
                buf.copy_(t.reshape(buf.shape))              # converts dtype (fp32 -> bf16 features) on the host
                out.append(buf)
        return out

    def _feed(self, batch_c3d, batch_maps, local=False):
        """-> device placeholders (asynchronous copies on the model's stream)"""
        c3d_h, gt_h = self._stage(batch_c3d, batch_maps, local=local)
        self.c3d_input.copy_(c3d_h, non_blocking=True)
        self.gt_gazemap.copy_(gt_h, non_blocking=True)

    def _flip_fed(self, clip_indices):
        """gaze_rnn.py:507-509 on the device placeholders: reverse the W axis of c3d_input and gt_gazemap for the
        chosen clips (global clip indices; this rank flips the ones that fell into its shard)."""
        lo = self.rank * self.local_batch
        mine = np.asarray([i - lo for i in clip_indices if lo <= i < lo + self.local_batch], dtype=np.int32)
        if mine.size == 0:
            return
        e = self.engine
        idx = torch.from_numpy(mine).to(e.device)
        s = torch.cuda.current_stream(e.device).cuda_stream if e.device.type == "cuda" else 0
        T, G, GH, GW = self.n_lstm_steps, e.G, self.gazemap_height, self.gazemap_width
        self._lib_handle().rgp_flip_w(self.c3d_input.data_ptr(), idx.data_ptr(), int(mine.size),
                                      T * self.dim_feature * G, G, s)
        self._lib_handle().rgp_flip_w(self.gt_gazemap.data_ptr(), idx.data_ptr(), int(mine.size), T * GH, GW, s)

    def _lib_handle(self):
        return self.engine.lib

    def _use_graph(self):
        e = self.engine
        return (bool(getattr(self.config, "cuda_graph", True)) and e.device.type == "cuda" and self.chunk_steps is None
                and self.model_name is not None)

    def _forward_fed(self):
        self.engine.forward(c3d=self.c3d_input, gt=self.gt_gazemap, loss_type=self.config.loss_type,
                            global_batch=self.batch_size)

    def _step_graph(self, train):
        """the captured step for this mode; re-captured when the optimiser hyper-parameters baked into the graph (clip,
        decay) change — the learning rate itself is a device scalar and needs no re-capture"""
        key = ("train", self.max_grad_norm, self.learning_rate_decay) if train else ("infer",)
        g = self._graphs.get(key)
        if g is None:
            if train and self.train_op is None:
                raise RuntimeError("this model was built without a train op")
            g = StepGraph(self.engine, self._forward_fed, train=train, allreduce=self._allreduce if train else None,
                          opt=self._opt_kwargs() if train else None)
            self._graphs[key] = g
        return g

    def run(self, train=False, sync=True):
        """One pass over the fed placeholders: forward + loss (+ BPTT + clip/Adam when train) — one CUDA-graph submission
        (config.cuda_graph), else eager launches.  sync=True returns the loss (a device->host read); sync=False leaves
        the read pending (``last_loss`` collects it) so that the host can prepare the next batch meanwhile."""
        if self.chunk_steps is not None:
            if train:
                raise RuntimeError("a chunked model cannot train")
            return self._run_chunked()
        if self.model_name is None:          # dense variant: live dropout while training (gaze_rnn.py:529 feeds 0.5)
            self.engine.keep_prob = float(getattr(self.config, "dropout_keep_prob_train", 0.5)) if train else 1.0
            self.engine.dropout_seed = (int(getattr(self.config, "seed", 0)) << 32) + int(self.current_step)
        if self._use_graph():
            if train:
                self.engine.set_lr0(self.initial_learning_rate * self._lr_scale)
            self._step_graph(train).replay()
            if train:
                self.global_step = self.engine.step
        else:
            self._forward_fed()
            if train:
                if self.train_op is None:
                    raise RuntimeError("this model was built without a train op")
                self.train_op()
        loss_dev = self.engine.loss
        if self.world_size > 1 and self._allreduce is not None:
            loss_dev = self.engine.loss.clone()
            self._allreduce(loss_dev)   # local losses are already scaled by 1/(B_global*T): the sum is the global loss
        if self._loss_event is None:
            self._last_loss, self._loss_pending = float(loss_dev.cpu()[0]), False
            return self._last_loss
        self._loss_host.copy_(loss_dev, non_blocking=True)
        self._loss_event.record(torch.cuda.current_stream(self.engine.device))
        self._loss_pending = True
        return self.last_loss if sync else None

    @property
    def last_loss(self):
        if self._loss_pending:
            self._loss_event.synchronize()
            self._last_loss, self._loss_pending = float(self._loss_host[0]), False
            # the persistent small-batch kernels never hang: a dependency wait that gives up is reported through a counter
            # and the pass is invalid - fail loudly where the loss is read (one 4-byte read, already synchronised)
            timed_out = getattr(self.engine, "seq_timed_out", None)
            if timed_out is not None and timed_out():
                raise RuntimeError("a persistent recurrence kernel timed out waiting for a neighbouring tile; results invalid")
        return self._last_loss

    @last_loss.setter
    def last_loss(self, v):
        self._last_loss, self._loss_pending = float(v), False

    def _run_chunked(self):
        """The clip in chunks of `chunk_steps` frames: the state left by chunk k (h_T, and the LSTM cell) is the initial
        state of chunk k+1 and `t0` selects the per-time-step BN parameters, so the result equals the one-pass model's."""
        e, Tc = self.engine, self.chunk_steps
        loss, nchunks = 0.0, self.n_lstm_steps // Tc
        for k in range(nchunks):
            t0 = k * Tc
            e.forward(c3d=self.c3d_input[:, t0:t0 + Tc].contiguous(), gt=self.gt_gazemap[:, t0:t0 + Tc].contiguous(),
                      loss_type=self.config.loss_type, global_batch=self.batch_size, carry_state=k > 0, t0=t0)
            self.predicted_gazemaps_logit[:, t0:t0 + Tc].copy_(e.logits)
            self._probs[:, t0:t0 + Tc].copy_(e.probs)
            loss += float(e.loss.cpu()[0]) / nchunks       # each chunk's loss is a mean over B*chunk_steps frames
            e.carry_last_state()
        e.reset_state()
        if self.world_size > 1 and self._allreduce is not None:
            t = torch.tensor([loss], dtype=torch.float32, device=e.device)
            self._allreduce(t)
            loss = float(t.cpu()[0])
        return loss

    def predict(self, batch_c3d, out=None, local=False):
        """The inference call the reference's extract_map.py:210-217 wishes the model had
        (`session.run([model.predicted_gazemaps], {model.c3d_input: ..})`): c3d [batch_size, T, 1024, 7, 7] host array
        (NumPy, or a CPU tensor — pinned memory is read in place by the copy engine; bf16 features are accepted when
        config.feature_dtype == 'bf16') -> gaze maps [local clips, T, GH, GW] (probabilities for xentropy / KLD, raw maps
        for l2).  `out`: optional pinned CPU tensor that receives the maps (returned as such, after synchronising);
        default: a fresh NumPy array."""
        c3d_h, _ = self._stage(batch_c3d, self._gt_host, local=True) if local else \
            self._stage(batch_c3d, np.zeros((self.batch_size,) + tuple(self._gt_host.shape[1:]), dtype=np.float32))
        self.c3d_input.copy_(c3d_h, non_blocking=True)
        self.run(train=False, sync=False)
        if out is None:
            return self.predicted_gazemaps.cpu().numpy()
        out.copy_(self.predicted_gazemaps, non_blocking=True)
        torch.cuda.current_stream(self.engine.device).synchronize()
        return out

    # ---------------------------------------------------------------- pipelined feeding (fit / benchmark loops)
    def _get_feeder(self):
        if self._feeder is None:
            from ..pipeline import PrefetchFeeder
            self._feeder = PrefetchFeeder([self.c3d_input, self.gt_gazemap], device=self.engine.device)
        return self._feeder

    def prefetch(self, batch_c3d, batch_maps, flip=None, local=False):
        """Start uploading a host batch on the copy stream while the current step runs (double-buffered: at most two
        batches in flight).  `acquire()` then makes it the fed batch.  Pinned inputs are read in place and must stay
        untouched until acquire(); pageable inputs are copied through one of two pinned staging sets first."""
        f = self._get_feeder()
        k = f.n_prefetched % f.slots
        if not hasattr(self, "_pf_stage"):
            pin = self._c3d_host.is_pinned()
            self._pf_stage = [(torch.zeros(tuple(self._c3d_host.shape), dtype=self._c3d_host.dtype, pin_memory=pin),
                               torch.zeros(tuple(self._gt_host.shape), dtype=torch.float32, pin_memory=pin))
                              for _ in range(f.slots)]
            self._pf_flip = [None] * f.slots
        f.free[k].synchronize()                      # the slot's previous upload has been consumed
        self._pf_flip[k] = flip
        f.prefetch(self._stage(batch_c3d, batch_maps, stage=self._pf_stage[k], local=local))

    def acquire(self):
        """the oldest prefetched batch becomes the content of the device placeholders (+ its flip, gaze_rnn.py:504-510)"""
        f = self._get_feeder()
        k = f.n_acquired % f.slots
        f.acquire()
        if self._pf_flip[k] is not None:
            self._flip_fed(self._pf_flip[k])

    def _host_batch(self, dataset, train_mode):
        """gaze_rnn.py:487-510: fetch, reshape, normalise the gaze maps, draw the flip (global numpy RNG, same draw)."""
        batch_images, batch_maps, batch_fixmaps, batch_c3d, batch_pupil, batch_clipnames = dataset.next_batch(self.batch_size)
        assert batch_images.dtype == np.float32
        batch_c3d = np.reshape(batch_c3d, [self.batch_size, -1, self.dim_feature, self.engine.G, self.engine.G])
        if self.config.loss_type in ('xentropy', 'KLD'):
            batch_maps = normalize_probability_map(batch_maps)
        flip = None
        if train_mode and self.config.use_flip_batch:        # gaze_rnn.py:504-510
            batch_size = len(batch_images)
            flip = np.random.choice(batch_size, batch_size // 2, replace=False)
            # images and fixation maps are not consumed by the step (gaze_grcn.py:173-376 never reads the frames):
            # their host arrays are flipped as in the reference; the two tensors the step DOES consume — C3D features
            # and gaze maps — are flipped on the device after the copy (rgp_flip_w), not in host memory
            if batch_images.shape[2] > 1:
                batch_images[flip, :] = batch_images[flip, :, :, ::-1, :]
            batch_fixmaps[flip, :] = batch_fixmaps[flip, :, :, ::-1]
        return batch_c3d, batch_maps, flip

    def single_step(self, train_mode=True, dataset=None, _prefetched=False, _prefetch_next=False):
        """gaze_rnn.py:483-565.  (_prefetched / _prefetch_next are fit()'s pipelining hooks: the batch of this step was
        uploaded during the previous one / the next training batch is prepared and uploaded while this step runs.)"""
        _start_time = time.time()
        if dataset is None:
            dataset = self.data_sets.train if train_mode else self.data_sets.valid
        if _prefetched:
            self.acquire()
        else:
            batch_c3d, batch_maps, flip = self._host_batch(dataset, train_mode)
            self._feed(batch_c3d, batch_maps)
            if flip is not None:
                self._flip_fed(flip)
        step = self.current_step + (1 if train_mode else 0)
        want_log = (not train_mode) or np.mod(step, self.config.steps_per_logprint) == 0
        self.run(train=train_mode, sync=False)               # asynchronous: the loss read stays pending
        if _prefetch_next:
            self.prefetch(*self._host_batch(self.data_sets.train, True))
        step = self.current_step
        if want_log:
            loss = self.last_loss                            # device -> host read (synchronises)
            _end_time = time.time()
            epoch = float(step * self.batch_size) / max(1, len(self.data_sets.train))
            (log.info if train_mode else log.warning)(
                " [%5s epoch %.1f / step %4d] %sbatch total-loss: %.5f, target-loss: %.5f "
                "(%.3f sec/batch, %.3f instances/sec) (lr=%.3g)",
                'train' if train_mode else 'val', epoch, step,
                (self.config.train_tag + ' | ') if self.config.train_tag else '', loss, loss,
                _end_time - _start_time, self.batch_size / (_end_time - _start_time), self.current_learning_rate)
        return step

    def generate(self, dataset, max_instances=50):
        """gaze_rnn.py:568-650: batched inference; per-frame lists under the reference's dict keys."""
        GH, GW = self.gazemap_height, self.gazemap_width
        pred, gts, fixs, images, names, c3ds = [], [], [], [], [], []
        n_instances = len(dataset)
        if max_instances is not None:
            n_instances = min(n_instances, max_instances)
        step_num = int(np.ceil(n_instances / float(self.batch_size)))
        assert step_num > 0
        for _ in range(step_num):
            batch_images, batch_maps, batch_fixmaps, batch_c3d, batch_pupil, batch_filename = dataset.next_batch(self.batch_size)
            batch_images = np.asarray(list(batch_images))
            assert batch_images.dtype == np.float32
            batch_c3d = np.reshape(batch_c3d, [self.batch_size, -1, self.dim_feature, self.engine.G, self.engine.G])
            if self.config.loss_type == 'xentropy':
                batch_maps = normalize_probability_map(batch_maps)
            self._feed(batch_c3d, batch_maps)
            self.run(train=False, sync=False)
            gazes = self._gather_maps(self.predicted_gazemaps)
            c3ds.extend(batch_c3d)
            pred.extend(gazes)
            gts.extend(batch_maps)
            fixs.extend(batch_fixmaps)
            images.extend(np.concatenate(batch_images))
            names.extend(batch_filename)
        pred = np.vstack(pred).reshape([-1, GH, GW])
        gts = np.vstack(gts).reshape([-1, GH, GW])
        c3ds = np.vstack(c3ds).reshape([-1, self.dim_feature, self.engine.G, self.engine.G])
        try:
            fixs = np.vstack(fixs)
        except ValueError:                  # per-clip fixation maps at differing resolutions (gaze_rnn.py:632-640)
            fixs = [fm[t] for fm in fixs for t in range(len(fm))]
            assert len(fixs) == len(pred)
        return {'pred_gazemap_list': pred, 'gt_gazemap_list': gts, 'images_list': images,
                'fixationmap_list': fixs, 'clipname_list': names, 'c3d_list': c3ds}

    def _gather_maps(self, maps_dev):
        """[local clips, T, GH, GW] device maps -> NumPy maps of ALL ranks' clips in rank order (the final
        gather of data-parallel inference): one all-gather of the device tensor into a preallocated device buffer
        (NCCL over NVLink on GPUs) and one device->host copy — no pickling."""
        if self.world_size == 1:
            return maps_dev.cpu().numpy()
        import torch.distributed as dist
        shape = (self.world_size * maps_dev.shape[0],) + tuple(maps_dev.shape[1:])     # rank-major concatenation
        if self._gather_buf is None or tuple(self._gather_buf.shape) != shape:
            self._gather_buf = torch.empty(shape, dtype=maps_dev.dtype, device=maps_dev.device)
        dist.all_gather_into_tensor(self._gather_buf, maps_dev.contiguous())
        return self._gather_buf.cpu().numpy()

    def evaluate(self, pred_gazemap_list, gt_gazemap_list, fixationmap_list, images_list, **_ignored):
        """gaze_rnn.py:653-674.  (`**_ignored` absorbs the extra keys generate() returns — the reference raises a
        TypeError there)"""
        assert len(pred_gazemap_list) == len(gt_gazemap_list) == len(fixationmap_list) == len(images_list), \
            "Length mismatch: %d %d %d %d" % (len(pred_gazemap_list), len(gt_gazemap_list), len(fixationmap_list),
                                              len(images_list))
        batch_scores = {}
        log.info('Evaluation on %d images', len(gt_gazemap_list))
        eng = EM.MetricsEngine(device=self.engine.device, lib=self._lib)
        for metric in EM.AVAILABLE_METRICS:
            batch_scores[metric] = EM.saliency_score(metric, pred_gazemap_list, gt_gazemap_list, fixationmap_list,
                                                     engine=eng)
            log.info('Saliency %s : %f', metric, batch_scores[metric])
        self.report_evaluate_summary(batch_scores)
        return batch_scores

    def generate_and_evaluate(self, dataset, max_instances=50):
        ret = self.generate(dataset, max_instances)
        scores = self.evaluate(**ret)
        return ret, scores
