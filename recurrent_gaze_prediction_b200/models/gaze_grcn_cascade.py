"""Two-level cascade GRCN — mirror of the reference's ``models/gaze_grcn_cascade.py``.

``GazePredictionGRCN`` (:152-479) keeps the reference's builder signature
``create_gazeprediction_network(frame_images, c3d_input, gt_gazemap, dropout_keep_prob, net=None)
-> (predicted_gazemaps, loss, image_summaries)``: unlike the other models this builder also owns the loss (summed l2,
:427-441) and emits the maps WITHOUT a softmax (:422-424).  The device work is :class:`CascadeEngine`.  Forward / loss /
generate / evaluate; the cascade's BPTT is not built (``config.trainable`` must be False).

The file as written cannot build its graph (top cell declared with 65 input channels, fed 64, :346-351,379).
``config.cascade_top_in`` = 64 (default, reading A: the cell takes what it is fed) or 65 (reading B: the commented-out
concat with the frame-saliency map, :366-374, is live; the map is the ``frm_sal`` placeholder because the ShallowNet
that would produce it from ``frame_images`` is outside the path).
"""
from __future__ import annotations

import numpy as np
import torch

from ..cascade_engine import CascadeEngine
from .gaze_grcn import GRU_RCN_Cell                                  # noqa: F401  (:48-146, the same cell class)
from .gaze_rnn import CONSTANTS, GRUModelConfig, GazePredictionGRU   # noqa: F401
from .model_util import init_parameters_cascade


class GazePredictionGRCN(GazePredictionGRU):
    model_name = "gaze_grcn_cascade"

    def build_model(self):
        """gaze_grcn_cascade.py:155-185."""
        c = self.config
        if bool(getattr(c, "trainable", True)):
            raise NotImplementedError("the cascade is built for inference / evaluation: set config.trainable = False")
        c.loss_type = 'l2'                                            # the builder's own loss (:427-441)
        B, T = self.local_batch, self.n_lstm_steps
        dev = torch.device(getattr(c, "device", "cuda"))
        dtype = 1 if getattr(c, "dtype", "bf16") == "bf16" else 0
        top_in = int(getattr(c, "cascade_top_in", 64))
        self.chunk_steps = None
        self.engine = CascadeEngine(B=B, T=T, Ch=getattr(c, "cascade_state_size", 256), Cin=self.dim_cnn_proj,
                                    Cfeat=c.dim_feature, dtype=dtype, device=dev, lib=self._lib, top_in=top_in)
        if (self.gazemap_height, self.gazemap_width) != (49, 49):
            raise ValueError("the cascade predicts 49x49 gaze maps")
        self.engine.load_tf_params(init_parameters_cascade(hidden=self.engine.Ch, top_in=top_in, dim_proj=self.dim_cnn_proj,
                                                           dim_feature=c.dim_feature, seed=getattr(c, "seed", 0)))
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        fdt = {'fp32': torch.float32, 'bf16': torch.bfloat16}[getattr(c, "feature_dtype", 'fp32')]
        self.frame_images = None          # consumed only by the ShallowNet branch, whose map never reaches the prediction
        self.frm_sal = z(B, T, 49, 49) if top_in == 65 else None
        self.c3d_input = torch.zeros(B, T, c.dim_feature, 7, 7, dtype=fdt, device=dev)
        self.gt_gazemap = z(B, T, 49, 49)
        self.dropout_keep_prob = 1.0      # live only while the reference trains (:401-402)
        self.predicted_gazemaps_logit = self.engine.logits
        self._probs = self.engine.logits
        self.predicted_gazemaps = self.engine.logits
        self.loss = self.engine.loss
        self._alloc_staging(dev, fdt)

    def _forward_fed(self):
        self.engine.forward(c3d=self.c3d_input, gt=self.gt_gazemap, loss_type='l2', global_batch=self.batch_size,
                            frm_sal=self.frm_sal)

    @staticmethod
    def create_gazeprediction_network(frame_images, c3d_input, gt_gazemap, dropout_keep_prob=None, net=None, params=None,
                                      dtype='fp32', seed=0, lib=None, frm_sal=None):
        """gaze_grcn_cascade.py:188-479 -> (predicted_gazemaps [B,T,49,49], loss, image_summaries=None) for c3d_input
        [B,T,1024,7,7] and gt_gazemap [B,T,49,49] (fp32 device tensors); fresh reference-initialised weights or `params`
        (scope-qualified reference names).  Fills net['c3d_embedded'-free keys]: 'rcn_outputs', 'rcn_upsampled_outputs',
        'predicted_gazemaps'."""
        if net is not None:
            assert isinstance(net, dict)
        if dropout_keep_prob is not None and float(dropout_keep_prob) != 1.0:
            raise NotImplementedError("dropout on fc1 (:401-402) is a training-time node; the cascade is forward-only here")
        B, T, F, G, G2 = c3d_input.shape
        assert B > 0 and T > 0 and G == 7 and G2 == 7
        assert tuple(gt_gazemap.shape) == (B, T, 49, 49)
        top_in = 64 if frm_sal is None else 65
        if params is not None:
            top_in = int(np.asarray(params["RCNGaze/GRU_Conv_Wz"]).shape[2])
        hidden = 256 if params is None else int(np.asarray(params["RCNBottom/GRU_Conv_Uz"]).shape[-1])
        dim_proj = 512 if params is None else int(np.asarray(params["proj_c3d_W"]).shape[1])
        eng = CascadeEngine(B=B, T=T, Ch=hidden, Cin=dim_proj, Cfeat=F, dtype=1 if dtype == 'bf16' else 0,
                            device=c3d_input.device, lib=lib, top_in=top_in)
        eng.load_tf_params(params if params is not None else
                           init_parameters_cascade(hidden=hidden, top_in=top_in, dim_proj=dim_proj, dim_feature=F, seed=seed))
        pred = eng.forward(c3d=c3d_input.contiguous(), gt=gt_gazemap.contiguous(), frm_sal=frm_sal)
        if net is not None:
            net['rcn_outputs'] = [eng.read_state(t + 1) for t in range(T)]
            net['rcn_upsampled_outputs'] = [eng.read_upsampled(t) for t in range(T)]
            net['predicted_gazemaps'] = pred
            net['engine'] = eng
        return pred, eng.loss, None
