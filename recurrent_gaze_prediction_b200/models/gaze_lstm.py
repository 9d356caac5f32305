"""ConvLSTM gaze model — mirror of the reference's ``models/gaze_lstm.py``.

``LSTM_RCN_Cell`` (:48-148) and ``GazePredictionLSTM`` (:154-353).  The cell arithmetic is the reference's AS WRITTEN
(SURVEY Q7): the candidate re-uses the W_hi convolution, the output gate peeps at the OLD cell, peepholes are full
[H,W,C] tensors, W_hc exists but is dead, state = concat([c, h], axis=3)."""
from __future__ import annotations

import numpy as np
import torch

from ..engine import GazeEngine
from .gaze_grcn import _create_network
from .gaze_rnn import CONSTANTS, GRUModelConfig, GazePredictionGRU   # noqa: F401
from .model_util import truncated_normal


class LSTM_RCN_Cell(object):
    """gaze_lstm.py:48-148.  ``state`` is [B,H,W,2*num_units] = concat([c, h], 3) (fp32 device tensor)."""

    def __init__(self, num_units, dim_feature, spatial_shape=[7, 7], kernel_spatial_shape=[3, 3],
                 dtype='fp32', device='cuda', seed=0, lib=None):
        if list(kernel_spatial_shape) != [3, 3] or spatial_shape[0] != spatial_shape[1]:
            raise ValueError("3x3 kernels on square grids only")
        self._num_units, self._dim_feature = num_units, dim_feature
        self._spatial_shape = list(spatial_shape)
        self._kernel_spatial_shape = list(kernel_spatial_shape)
        self._dtype, self._device, self._lib = dtype, device, lib
        rng = np.random.RandomState(seed)
        G = spatial_shape[0]
        self.variables = {}
        for g in "ifco":                               # creation order of gaze_lstm.py:64-88
            self.variables["ConvLSTM_Wx" + g] = truncated_normal(rng, (3, 3, dim_feature, num_units), 1e-4)
            self.variables["ConvLSTM_Wh" + g] = truncated_normal(rng, (3, 3, num_units, num_units), 1e-4)
            if g != "c":
                self.variables["ConvLSTM_Wc" + g] = truncated_normal(rng, (G, G, num_units), 1e-4)
        self._engine = None

    @property
    def input_size(self):
        return self._dim_feature

    @property
    def output_size(self):
        return self._num_units

    @property
    def state_size(self):
        return 2 * self._num_units

    def load_variables(self, variables):
        self.variables.update({k: np.asarray(v) for k, v in variables.items()})
        self._engine = None

    def _get_engine(self, B):
        if self._engine is None or self._engine.B != B:
            G = self._spatial_shape[0]
            e = GazeEngine("gaze_lstm", B=B, T=1, G=G, Ch=self._num_units, Cin=self._dim_feature, Cfeat=64,
                           dtype=1 if self._dtype == 'bf16' else 0, device=self._device, lib=self._lib, head="fc")
            p = dict(self.variables)
            p["out_W"] = np.zeros((self._num_units, 1))
            p["out_b"] = np.zeros((1,))
            e.load_tf_params(p)
            self._engine = e
        return self._engine

    def __call__(self, inputs, state, scope=None):
        """gaze_lstm.py:103-133 -> (new_h, concat([new_c, new_h], 3))."""
        B, H, W, C = inputs.shape
        assert [H, W] == self._spatial_shape, "spatial shape mismatch"
        n = self._num_units
        assert C == self._dim_feature and tuple(state.shape) == (B, H, W, 2 * n)
        e = self._get_engine(B)
        e.set_initial_state(state[..., n:].contiguous(), state[..., :n].contiguous())
        e.forward(emb=inputs.reshape(B, 1, H, W, C).contiguous(), carry_state=True)
        c, h = e.read_state(1)
        return h, torch.cat([c, h], dim=3)

    def zero_state(self, batch_size, dtype=torch.float32):
        """gaze_lstm.py:136-148."""
        H, W = self._spatial_shape
        return torch.zeros(batch_size, H, W, 2 * self._num_units, dtype=dtype, device=self._device)


class GazePredictionLSTM(GazePredictionGRU):
    """gaze_lstm.py:154-353 (49x49 maps; no batch-norm scale in front of the head)."""
    model_name = "gaze_lstm"

    @staticmethod
    def create_gazeprediction_network(frame_images, c3d_input, dropout_keep_prob=1.0, net=None, params=None,
                                      dtype='fp32', rnn_state_size=128, seed=0, lib=None):
        """gaze_lstm.py:177-353 -> logits [B,T,49,49]."""
        return _create_network("gaze_lstm", c3d_input, net, params, dtype, rnn_state_size, seed, lib)
