"""Host helpers of the model classes (reference: ``models/model_util.py``) and the reference's parameter
initialisers.  NumPy only — these run once per batch / once per model on small host arrays; the device work
(softmax, loss) lives in ``csrc/ops.cu``."""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np


def normalize_probability_map(t):
    """models/model_util.py:40-58: every [.., H, W] map of a [B,H,W] or [B,T,H,W] batch is divided by its sum."""
    t = np.array(t, copy=True)
    if t.ndim not in (3, 4):
        raise ValueError('Unsupported shape : {}'.format(t.shape))
    s = t.reshape(t.shape[:-2] + (-1,)).sum(axis=-1)
    t /= s[..., None, None]
    return t


def truncated_normal(rng, shape, stddev):
    """tf.truncated_normal_initializer: samples beyond 2 sigma are re-drawn (gaze_grcn.py:64-81)."""
    x = rng.standard_normal(size=shape)
    bad = np.abs(x) > 2.0
    while bad.any():
        x[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(x) > 2.0
    return x * stddev


def xavier_uniform_conv2d(rng, shape):
    """tf.contrib.layers.xavier_initializer_conv2d(uniform=True) on a [kH,kW,A,B] kernel (gaze_grcn.py:292-310)."""
    rf = shape[0] * shape[1]
    lim = math.sqrt(6.0 / (rf * shape[2] + rf * shape[3]))
    return rng.uniform(-lim, lim, size=shape)


def init_parameters(model="gaze_grcn", hidden=128, T=35, grid=7, dim_proj=512, dim_feature=1024, seed=0,
                    head=None, stddev=1e-4):
    """Fresh parameters under the reference's variable names and TF layouts, created in the reference's
    variable-creation order: proj_c3d_W/b U(-0.1,0.1) (gaze_grcn.py:234-237); cell kernels truncated-normal
    sigma=1e-4 in the order of GRU_RCN_Cell.__init__ (:64-81) / LSTM_RCN_Cell.__init__ (gaze_lstm.py:64-88);
    Upsampling/weight1..3 Xavier-uniform (:292-310); out_W/out_b U(-0.1,0.1) (:311-314); one (gamma_t, beta_t) = (1, 0)
    pair per time step for the frozen batch-norm (:325)."""
    rng = np.random.RandomState(seed)
    if head is None:
        head = "fc" if model == "gaze_grcn77" else "deconv"
    p = OrderedDict()
    p["proj_c3d_W"] = rng.uniform(-0.1, 0.1, size=(dim_feature, dim_proj))
    p["proj_c3d_b"] = rng.uniform(-0.1, 0.1, size=(dim_proj,))
    if model == "gaze_lstm":
        for g in "ifco":
            p["ConvLSTM_Wx" + g] = truncated_normal(rng, (3, 3, dim_proj, hidden), stddev)
            p["ConvLSTM_Wh" + g] = truncated_normal(rng, (3, 3, hidden, hidden), stddev)
            if g != "c":
                p["ConvLSTM_Wc" + g] = truncated_normal(rng, (grid, grid, hidden), stddev)
    else:
        for nm, cin in (("Wz", dim_proj), ("Uz", hidden), ("Wr", dim_proj), ("Ur", hidden), ("W", dim_proj),
                        ("U", hidden)):
            p["GRU_Conv_" + nm] = truncated_normal(rng, (3, 3, cin, hidden), stddev)
    if head == "deconv":
        p["Upsampling/weight1"] = xavier_uniform_conv2d(rng, (5, 5, 64, hidden))
        p["Upsampling/weight2"] = xavier_uniform_conv2d(rng, (5, 5, 32, 64))
        p["Upsampling/weight3"] = xavier_uniform_conv2d(rng, (7, 7, 12, 32))
        p["out_W"] = rng.uniform(-0.1, 0.1, size=(12, 1))
    else:
        p["out_W"] = rng.uniform(-0.1, 0.1, size=(hidden, 1))
    p["out_b"] = rng.uniform(-0.1, 0.1, size=(1,))
    if model == "gaze_grcn" and head == "deconv":
        p["bn_gamma"] = np.ones((T, hidden))
        p["bn_beta"] = np.zeros((T, hidden))
    return p


def init_parameters_dense(dim_proj=32, grid=7, dim_feature=1024, GH=49, GW=49, seed=0):
    """Variables of the dense GRU variant (gaze_rnn.py:294-321) under their TF names, in creation order:
    proj_c3d_W/b U(-0.1,0.1); RNN/gru_cell/{gates,candidate}/{kernel,bias} (orthogonal kernels — the reference names the
    initializer "orthogonal" at :315 — biases 1 / 0 as tf.nn.rnn_cell.GRUCell creates them); proj_out_W U(-0.1,0.1),
    proj_out_b zeros."""
    rng = np.random.RandomState(seed)
    I = grid * grid * dim_proj
    H = I + grid * grid                                         # gaze_rnn.py:243-246: 7*7*32 + 7*7 = 1617
    p = OrderedDict()
    p["proj_c3d_W"] = rng.uniform(-0.1, 0.1, size=(dim_feature, dim_proj))
    p["proj_c3d_b"] = rng.uniform(-0.1, 0.1, size=(dim_proj,))

    def orth(rows, cols):
        q, r = np.linalg.qr(rng.standard_normal((max(rows, cols), min(rows, cols))))
        q = q * np.sign(np.diag(r))
        return q if rows >= cols else q.T

    p["RNN/gru_cell/gates/kernel"] = orth(I + H, 2 * H)
    p["RNN/gru_cell/gates/bias"] = np.ones(2 * H)
    p["RNN/gru_cell/candidate/kernel"] = orth(I + H, H)
    p["RNN/gru_cell/candidate/bias"] = np.zeros(H)
    p["proj_out_W"] = rng.uniform(-0.1, 0.1, size=(H, GH * GW))
    p["proj_out_b"] = np.zeros(GH * GW)
    return p


def xavier_uniform(rng, fan_in, fan_out):
    """tf.contrib.layers.xavier_initializer(uniform=True) on a [fan_in, fan_out] matrix (gaze_grcn_cascade.py:390,405)."""
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=(fan_in, fan_out))


def init_parameters_cascade(hidden=256, top_in=64, units=3, dim_proj=512, dim_feature=1024, seed=0, stddev=1e-4):
    """Variables of the cascade (gaze_grcn_cascade.py) under scope-qualified TF names, in creation order: proj_c3d_W/b
    U(-0.1,0.1) (:268-269); RCNBottom cell, 3x3, truncated normal sigma=1e-4 (:52-81 via :287); Upsampling/weight
    [11,11,64,hidden] Xavier-uniform (:317-321); RCNGaze cell, 5x5, `top_in` input channels, 3 units (:346-351: declared
    65, fed 64 - see cascade_engine.py); LastProjection/fc1 [7203,4802] and fc2 [2401,4802] Xavier-uniform with zero
    biases (:388-411)."""
    rng = np.random.RandomState(seed)
    p = OrderedDict()
    p["proj_c3d_W"] = rng.uniform(-0.1, 0.1, size=(dim_feature, dim_proj))
    p["proj_c3d_b"] = rng.uniform(-0.1, 0.1, size=(dim_proj,))
    for scope, k, cin, n in (("RCNBottom", 3, dim_proj, hidden), (None, 0, 0, 0), ("RCNGaze", 5, top_in, units)):
        if scope is None:
            p["Upsampling/weight"] = xavier_uniform_conv2d(rng, (11, 11, 64, hidden))
            continue
        for nm, ci in (("Wz", cin), ("Uz", n), ("Wr", cin), ("Ur", n), ("W", cin), ("U", n)):
            p[scope + "/GRU_Conv_" + nm] = truncated_normal(rng, (k, k, ci, n), stddev)
    p["LastProjection/fc1/weights"] = xavier_uniform(rng, 49 * 49 * units, 4802)
    p["LastProjection/fc1/bias"] = np.zeros(4802)
    p["LastProjection/fc2/weights"] = xavier_uniform(rng, 2401, 4802)
    p["LastProjection/fc2/bias"] = np.zeros(4802)
    return p
