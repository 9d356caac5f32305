"""GRCN with the 7x7 fully-connected read-out — mirror of the reference's ``models/gaze_grcn77.py``
(``GazePredictionGRCN`` :52-218: same ConvGRU, ``logit = h . out_W[Ch,1] + out_b`` per position :206-211)."""
from __future__ import annotations

from .gaze_grcn import GRU_RCN_Cell, _create_network            # noqa: F401
from .gaze_rnn import GRUModelConfig, GazePredictionGRU, _E

CONSTANTS = _E(image_width=98, image_height=98, gazemap_width=7, gazemap_height=7,
               saliencymap_width=49, saliencymap_height=49)       # gaze_grcn77.py:39-45


class GazePredictionGRCN(GazePredictionGRU):
    model_name = "gaze_grcn77"

    def _default_gazemap_hw(self):
        g = getattr(self.config, "grid", 7)
        return g, g

    @staticmethod
    def create_gazeprediction_network(frame_images, c3d_input, dropout_keep_prob=1.0, net=None, params=None,
                                      dtype='fp32', rnn_state_size=128, seed=0, lib=None):
        """gaze_grcn77.py:76-218 -> logits [B,T,7,7]."""
        return _create_network("gaze_grcn77", c3d_input, net, params, dtype, rnn_state_size, seed, lib)
