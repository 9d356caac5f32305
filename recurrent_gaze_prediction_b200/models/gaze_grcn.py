"""GRCN (convolutional GRU) gaze model — mirror of the reference's ``models/gaze_grcn.py``.

``GRU_RCN_Cell`` (:48-146) and ``GazePredictionGRCN`` (:152-376) keep their reference signatures.  The cell's
``__call__`` runs ONE recurrence step through the same tap-GEMM kernels the unrolled network uses (input-half 3x3
conv, U_z/U_r GEMM with the sigmoid / r.h epilogue, U GEMM with the tanh / state-update epilogue); the model class
runs all T frames: projection -> input-half conv of all frames at once -> T recurrent steps -> composed
transposed-conv head -> spatial softmax + loss.
"""
from __future__ import annotations

import numpy as np
import torch

from ..engine import GazeEngine
from .gaze_rnn import CONSTANTS, GRUModelConfig, GazePredictionGRU   # noqa: F401  (re-exported like the reference)
from .model_util import init_parameters, truncated_normal


class GRU_RCN_Cell(object):
    """gaze_grcn.py:48-146.  ``inputs`` [B,H,W,dim_feature] and ``state`` [B,H,W,num_units] are fp32 device tensors."""

    def __init__(self, num_units, dim_feature, spatial_shape=[7, 7], kernel_spatial_shape=[3, 3],
                 dtype='fp32', device='cuda', seed=0, lib=None):
        if list(kernel_spatial_shape) != [3, 3]:
            raise ValueError("only 3x3 kernels are implemented (the reference never uses another size)")
        if spatial_shape[0] != spatial_shape[1]:
            raise ValueError("square grids only")
        self._num_units = num_units
        self._dim_feature = dim_feature
        self._spatial_shape = list(spatial_shape)
        self._kernel_spatial_shape = list(kernel_spatial_shape)
        self._dtype, self._device, self._lib = dtype, device, lib
        rng = np.random.RandomState(seed)
        # gaze_grcn.py:64-81: truncated normal, sigma = 1e-4, no biases, this creation order
        self.variables = {}
        for nm, cin in (("Wz", dim_feature), ("Uz", num_units), ("Wr", dim_feature), ("Ur", num_units),
                        ("W", dim_feature), ("U", num_units)):
            self.variables["GRU_Conv_" + nm] = truncated_normal(rng, (3, 3, cin, num_units), 1e-4)
        self._engine = None

    @property
    def input_size(self):
        return self._dim_feature

    @property
    def output_size(self):
        return self._num_units

    @property
    def state_size(self):
        return self._num_units

    def _get_engine(self, B):
        if self._engine is None or self._engine.B != B:
            G = self._spatial_shape[0]
            e = GazeEngine("gaze_grcn77", B=B, T=1, G=G, Ch=self._num_units, Cin=self._dim_feature, Cfeat=64,
                           dtype=1 if self._dtype == 'bf16' else 0, device=self._device, lib=self._lib)
            p = dict(self.variables)
            p["out_W"] = np.zeros((self._num_units, 1))
            p["out_b"] = np.zeros((1,))
            e.load_tf_params(p)
            self._engine = e
        return self._engine

    def load_variables(self, variables):
        self.variables.update({k: np.asarray(v) for k, v in variables.items()})
        self._engine = None

    def __call__(self, inputs, state, scope=None):
        """gaze_grcn.py:95-129 -> (output, new_state), both [B,H,W,num_units]."""
        B, H, W, C = inputs.shape
        assert [H, W] == self._spatial_shape, "spatial shape mismatch"        # gaze_grcn.py:97-100
        assert C == self._dim_feature and tuple(state.shape) == (B, H, W, self._num_units)
        e = self._get_engine(B)
        e.set_initial_state(state)
        e.forward(emb=inputs.reshape(B, 1, H, W, C).contiguous(), carry_state=True)
        h = e.read_state(1)
        return h, h

    def zero_state(self, batch_size, dtype=torch.float32):
        """gaze_grcn.py:132-146."""
        H, W = self._spatial_shape
        return torch.zeros(batch_size, H, W, self._num_units, dtype=dtype, device=self._device)


class GazePredictionGRCN(GazePredictionGRU):
    """gaze_grcn.py:152-376 (49x49 gaze maps through the transposed-conv head)."""
    model_name = "gaze_grcn"

    @staticmethod
    def create_gazeprediction_network(frame_images, c3d_input, dropout_keep_prob=1.0, net=None, params=None,
                                      dtype='fp32', rnn_state_size=128, seed=0, lib=None):
        """gaze_grcn.py:173-376: logits [B,T,49,49] of freshly initialised (or given ``params``, reference variable
        names) weights for c3d_input [B,T,1024,7,7] (fp32 device tensor).  ``frame_images`` is accepted and ignored, as
        in the reference graph.  Fills net['c3d_embedded'], net['rcn_outputs'], net['predicted_gazemaps']."""
        return _create_network("gaze_grcn", c3d_input, net, params, dtype, rnn_state_size, seed, lib)


def _create_network(model, c3d_input, net, params, dtype, rnn_state_size, seed, lib):
    if net is not None:
        assert isinstance(net, dict)
    B, T, F, G, G2 = c3d_input.shape
    assert B > 0 and T > 0 and G == G2
    dim_cnn_proj = 512                                                    # gaze_grcn.py:204
    if params is not None and "proj_c3d_W" in params:
        dim_cnn_proj = np.asarray(params["proj_c3d_W"]).shape[1]
    eng = GazeEngine(model, B=B, T=T, G=G, Ch=rnn_state_size, Cin=dim_cnn_proj, Cfeat=F,
                     dtype=1 if dtype == 'bf16' else 0, device=c3d_input.device, lib=lib)
    eng.load_tf_params(params if params is not None else
                       init_parameters(model, hidden=rnn_state_size, T=T, grid=G, dim_proj=dim_cnn_proj, dim_feature=F,
                                       seed=seed))
    logits = eng.forward(c3d=c3d_input.contiguous())
    if net is not None:
        n = B * eng.S
        xp = eng.Xp[:T * n].view(T, B, eng.Wp, eng.Wp, eng.Cin)[:, :, 1:, 1:, :]
        net['c3d_embedded'] = xp.permute(1, 0, 2, 3, 4).float()                       # [B,T,7,7,512]
        net['rcn_outputs'] = [eng.read_state(t + 1)[1] if eng.lstm else eng.read_state(t + 1) for t in range(T)]
        net['predicted_gazemaps'] = logits
        net['engine'] = eng
    return logits
