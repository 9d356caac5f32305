"""Model base class and configuration — mirror of the reference's ``models/base.py``.

Same surface (``BaseModelConfig`` :19-86, ``ModelBase`` :96-404: ``reload_checkpoint``, ``load_model_checkpoint``,
``load_model_from_checkpoint_file``, ``save_model_checkpoint``, ``fit``, ``report_evaluate_summary``,
``current_step``, ``current_learning_rate``, ``decay_learning_rate``), with the TensorFlow session replaced by a
:class:`~recurrent_gaze_prediction_b200.engine.GazeEngine` that owns the parameters and optimiser state on the GPU.
Checkpoints are ``.npz`` files holding the variables under the reference's TF names (TF layouts), the Adam moments
and the global step; ``<train_dir>/model/checkpoint`` names the latest one like ``tf.train.get_checkpoint_state``.
"""
from __future__ import annotations

import json
import logging
import os

import numpy as np

log = logging.getLogger("rgp")


class BaseModelConfig(object):
    """models/base.py:19-50 — same attribute names and defaults."""

    def __init__(self, train_dir=None, max_steps=100000, steps_per_checkpoint=1000, steps_per_validation=100,
                 steps_per_evaluation=2000, steps_per_logprint=5, verbose_level=2, learning_rate_decay=0.80,
                 optimization_method='adam', initial_learning_rate=0.003):
        self.train_dir = train_dir
        self.train_tag = ''
        self.max_steps = max_steps
        self.steps_per_checkpoint = steps_per_checkpoint
        self.steps_per_validation = steps_per_validation
        self.steps_per_evaluation = steps_per_evaluation
        self.steps_per_logprint = steps_per_logprint
        self.verbose_level = int(verbose_level)
        self.learning_rate_decay = learning_rate_decay
        self.optimization_method = optimization_method
        self.initial_learning_rate = initial_learning_rate
        self.max_grad_norm = 10.0

    def __repr__(self):
        return 'ModelConfig{' + ', '.join("%s : %s" % (k, v) for k, v in sorted(vars(self).items())) + '}'

    def dump(self, fp):
        """models/base.py:60-72: JSON of the attribute dict."""
        json.dump(vars(self), fp, indent=2, sort_keys=True)
        fp.write("\n")

    @classmethod
    def load(cls, fp):
        """models/base.py:74-85."""
        c = cls()
        for k, v in json.load(fp).items():
            setattr(c, k, v)
        return c


class ModelBase(object):
    """models/base.py:96-404.  Subclasses set ``self.engine`` (GazeEngine) in ``build_model``."""

    def __init__(self, config):
        self.config = config
        self.train_dir = config.train_dir
        self.initial_learning_rate = config.initial_learning_rate
        self.learning_rate_decay = config.learning_rate_decay
        self.max_grad_norm = config.max_grad_norm
        self.evaluation_history = []
        self._lr_scale = 1.0          # decay_learning_rate multiplies this (base.py:398-404)

    # ---------------------------------------------------------------- checkpoints (base.py:188-253)
    def reload_checkpoint(self):
        if self.train_dir and self.load_model_checkpoint(self.train_dir):
            log.info(" [Checkpoint] Successfully loaded model.")
            log.info("  Learning Rate = %.6f" % self.current_learning_rate)
            log.info("   Current Step = %d" % self.current_step)
            return True
        log.error(" [Checkpoint] Failed to load model !! (starting from scratch)")
        return False

    def load_model_checkpoint(self, checkpoint_dir):
        checkpoint_dir = os.path.join(checkpoint_dir, "model")
        state = os.path.join(checkpoint_dir, "checkpoint")
        if not os.path.isfile(state):
            return False
        with open(state) as fp:
            name = fp.read().strip()
        path = name if os.path.isabs(name) else os.path.join(checkpoint_dir, name)
        if not os.path.isfile(path):
            return False
        self.load_model_from_checkpoint_file(path)
        return True

    def load_model_from_checkpoint_file(self, checkpoint_path):
        import torch
        with np.load(checkpoint_path) as z:
            names = [k for k in z.files if not k.startswith("__")]
            self.engine.load_tf_params({k: z[k] for k in names})
            if self.engine.train and "__adam_m__" in z.files:
                self.engine.adam_m.copy_(torch.from_numpy(z["__adam_m__"]))
                self.engine.adam_v.copy_(torch.from_numpy(z["__adam_v__"]))
            if "__global_step__" in z.files and self.engine.train:
                self.engine.step = int(z["__global_step__"])
            if "__lr_scale__" in z.files:          # decay_learning_rate()'s accumulated factor (base.py:398-404)
                self._lr_scale = float(z["__lr_scale__"])
        log.info(" [Checkpoint] Successfully loaded from %s", checkpoint_path)

    def save_model_checkpoint(self, checkpoint_dir):
        """base.py:240-253.  Data-parallel ranks hold identical parameters: rank 0 alone writes (to a temporary file
        that is atomically renamed into place, so a reader never sees a partial archive), then all ranks meet at a
        barrier so nobody reloads before the file exists."""
        model_name = type(self).__name__ or "Model"
        checkpoint_dir = os.path.join(checkpoint_dir, "model")
        step = self.current_step
        fname = "%s-%d.npz" % (model_name, step)
        checkpoint_path = os.path.join(checkpoint_dir, fname)
        rank, world = int(getattr(self, "rank", 0)), int(getattr(self, "world_size", 1))
        if rank == 0:
            os.makedirs(checkpoint_dir, exist_ok=True)
            variables = self.engine.op_to_tf(self.engine.params.cpu().numpy())
            extra = {"__global_step__": np.int64(step), "__lr_scale__": np.float64(self._lr_scale)}
            if self.engine.train:
                extra["__adam_m__"] = self.engine.adam_m.cpu().numpy()
                extra["__adam_v__"] = self.engine.adam_v.cpu().numpy()
            tmp = checkpoint_path + ".tmp-%d.npz" % os.getpid()
            np.savez(tmp, **variables, **extra)
            os.replace(tmp, checkpoint_path)
            state_tmp = os.path.join(checkpoint_dir, "checkpoint.tmp-%d" % os.getpid())
            with open(state_tmp, "w") as fp:
                fp.write(fname + "\n")
            os.replace(state_tmp, os.path.join(checkpoint_dir, "checkpoint"))
            log.info(" [Checkpoint] Saved checkpoints into %s !" % checkpoint_path)
        if world > 1:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                dist.barrier()
        return checkpoint_path

    # ---------------------------------------------------------------- training loop (base.py:330-359)
    def fit(self):
        """base.py:330-359.  With config.prefetch the NEXT training batch is fetched, normalised and uploaded while the
        current step runs on the GPU (the step is submitted asynchronously; the loss is read only when it is logged).
        The order of every draw from the global numpy RNG is the reference's: a step that is followed by an evaluation
        (which draws jitter / negative samples) does not prefetch across it."""
        log.info("Initial Learning Rate = %.6f" % self.current_learning_rate)
        assert self.current_learning_rate > 0
        self.reload_checkpoint()
        c = self.config
        pipelined = bool(getattr(c, "prefetch", False)) and hasattr(self, "prefetch") and \
            getattr(self, "chunk_steps", None) is None and getattr(self.engine, "device", None) is not None and \
            self.engine.device.type == "cuda"
        step = 0
        have_next = False
        while step <= c.max_steps:
            nxt = self.current_step + 1
            crosses_eval = np.mod(nxt, c.steps_per_evaluation) == 0
            want_next = pipelined and nxt <= c.max_steps and not crosses_eval
            step = self.single_step(train_mode=True, _prefetched=have_next, _prefetch_next=want_next) if pipelined \
                else self.single_step(train_mode=True)
            have_next = want_next
            assert step > 0
            if np.mod(step, c.steps_per_checkpoint) == 0 and self.train_dir:
                self.save_model_checkpoint(self.train_dir)
            if np.mod(step, c.steps_per_validation) == 0:
                _ = self.single_step(train_mode=False)
            if np.mod(step, c.steps_per_evaluation) == 0:
                self.generate_and_evaluate(self.data_sets.valid)

    def report_evaluate_summary(self, batch_scores):
        """models/base.py:362-386 wrote TF summaries; here the scores are kept in ``evaluation_history``."""
        self.evaluation_history.append((self.current_step, dict(batch_scores)))

    @property
    def current_step(self):
        return int(getattr(self.engine, "step", 0))

    @property
    def current_learning_rate(self):
        return float(self._learning_rate_at(self.current_step))

    def _learning_rate_at(self, step):
        return self.initial_learning_rate * self._lr_scale

    def decay_learning_rate(self, decay_factor):
        self._lr_scale *= decay_factor
        return self.current_learning_rate
