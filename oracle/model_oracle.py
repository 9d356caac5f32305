"""CPU oracle (torch, CPU) for the RGP model path — TEST INFRASTRUCTURE ONLY.

A literal restatement of the reference's TF1 graph for the recurrent gaze
prediction path, keeping the reference's op structure (T-unrolled Python loop,
one conv per gate per frame, NHWC tensors, separate softmax and loss).  Run in
``torch.float64`` it is the parity truth; run in ``torch.float32`` it models
TF's fp32 rounding and is the timed "CPU restatement (torch/oneDNN), not
TensorFlow" baseline of bench.py.  BPTT gradients come from torch autograd on
this forward (the analogue of ``tf.gradients``, reference ``models/base.py:278-281``).

Every function cites the reference lines it follows (paths relative to
``/root/reference``).  Parity unpinned at the TF-library boundary: see
``oracle/__init__.py``.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-3  # tf.layers.batch_normalization default epsilon [TF-lib]


# ----------------------------------------------------------------------------
# TF op semantics restated on NHWC tensors
# ----------------------------------------------------------------------------
def conv2d_same(x, w):
    """tf.nn.conv2d(x, w, [1,1,1,1], 'SAME'): x [B,H,W,Cin], w [kH,kW,Cin,Cout],
    cross-correlation, zero padding (k-1)/2 (odd k). Used at models/gaze_grcn.py:108-125."""
    kh, kw = w.shape[0], w.shape[1]
    y = F.conv2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), padding=(kh // 2, kw // 2))
    return y.permute(0, 2, 3, 1)


def conv2d_transpose(x, w, stride, padding):
    """tf.nn.conv2d_transpose: x [B,H,W,Cin], w [kH,kW,Cout,Cin] (note Cout before Cin).
    VALID: out = (in-1)*s + k, out[b, i*s+ki, j*s+kj, co] += x[b,i,j,ci] * w[ki,kj,co,ci].
    SAME (stride 1, odd k): the VALID result cropped by (k-1)/2 on each side.
    Used at models/gaze_grcn.py:326-358."""
    kh = w.shape[0]
    pad = 0 if padding == "VALID" else (kh - 1) // 2
    if padding == "SAME":
        assert stride == 1
    y = F.conv_transpose2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), stride=stride, padding=pad)
    return y.permute(0, 2, 3, 1)


# ----------------------------------------------------------------------------
# Cells
# ----------------------------------------------------------------------------
def gru_rcn_cell(x, h, p):
    """GRU_RCN_Cell.__call__ — models/gaze_grcn.py:95-129.
    u(z) = sigmoid(Wz*x + Uz*h); r = sigmoid(Wr*x + Ur*h);
    c = tanh(W*x + U*(r.h)); h' = u.h + (1-u).c  (u=1 keeps state). No biases."""
    Wzconv = conv2d_same(x, p["GRU_Conv_Wz"])
    Uzconv = conv2d_same(h, p["GRU_Conv_Uz"])
    Wrconv = conv2d_same(x, p["GRU_Conv_Wr"])
    Urconv = conv2d_same(h, p["GRU_Conv_Ur"])
    u = torch.sigmoid(Wzconv + Uzconv)
    r = torch.sigmoid(Wrconv + Urconv)
    Wconv = conv2d_same(x, p["GRU_Conv_W"])
    Uconv = conv2d_same(r * h, p["GRU_Conv_U"])
    c = torch.tanh(Wconv + Uconv)
    new_h = u * h + (1 - u) * c
    return new_h, new_h


def lstm_rcn_cell(x, state, p):
    """LSTM_RCN_Cell.__call__ — models/gaze_lstm.py:103-133, bugs preserved:
    the candidate uses W_hi (not W_hc, which is dead); the output gate peeps at
    the OLD c; peepholes are full [H,W,Ch] tensors; state = concat([c,h], axis=3)."""
    ch = p["ConvLSTM_Wci"].shape[-1]
    c, h = state[..., :ch], state[..., ch:]
    i = torch.sigmoid(conv2d_same(x, p["ConvLSTM_Wxi"]) + conv2d_same(h, p["ConvLSTM_Whi"]) + p["ConvLSTM_Wci"] * c)
    f = torch.sigmoid(conv2d_same(x, p["ConvLSTM_Wxf"]) + conv2d_same(h, p["ConvLSTM_Whf"]) + p["ConvLSTM_Wcf"] * c)
    new_c = f * c + i * torch.tanh(conv2d_same(x, p["ConvLSTM_Wxc"]) + conv2d_same(h, p["ConvLSTM_Whi"]))
    o = torch.sigmoid(conv2d_same(x, p["ConvLSTM_Wxo"]) + conv2d_same(h, p["ConvLSTM_Who"]) + p["ConvLSTM_Wco"] * c)
    new_h = torch.tanh(new_c) * o
    return new_h, torch.cat([new_c, new_h], dim=3)


# ----------------------------------------------------------------------------
# Networks
# ----------------------------------------------------------------------------
def project_c3d(c3d, p):
    """models/gaze_grcn.py:223-254: [B,T,1024,7,7] -> transpose to NHWC ->
    xw_plus_b with proj_c3d_W [1024,512] -> [B,T,7,7,512]. Dropout is identity."""
    B, T = c3d.shape[:2]
    x = c3d.permute(0, 1, 3, 4, 2)
    G1, G2, K = x.shape[2], x.shape[3], x.shape[4]
    emb = x.reshape(-1, K) @ p["proj_c3d_W"] + p["proj_c3d_b"]
    return emb.reshape(B, T, G1, G2, -1)


def unpool_head(h, p, t, use_bn):
    """models/gaze_grcn.py:318-366 for one frame batch h [B,7,7,Ch]:
    BN(inference, per-timestep gamma/beta) -> 5x5 s3 VALID -> 5x5 s2 VALID ->
    7x7 s1 SAME -> xw_plus_b(12->1) -> [B,49,49]. No nonlinearity in between."""
    if use_bn:
        h = p["bn_gamma"][t] * h / math.sqrt(1.0 + BN_EPS) + p["bn_beta"][t]
    y = conv2d_transpose(h, p["Upsampling/weight1"], 3, "VALID")
    y = conv2d_transpose(y, p["Upsampling/weight2"], 2, "VALID")
    y = conv2d_transpose(y, p["Upsampling/weight3"], 1, "SAME")
    B, GH, GW, C = y.shape
    out = y.reshape(-1, C) @ p["out_W"] + p["out_b"]
    return out.reshape(B, GH, GW)


def recurrent_network(model, p, c3d=None, emb=None, net=None):
    """create_gazeprediction_network of gaze_grcn.py:173-376 / gaze_grcn77.py:94-218 /
    gaze_lstm.py:177-353.  `c3d` [B,T,1024,7,7] (reference entry) or `emb`
    [B,T,G,G,512] (BASELINE 'L1' features entering after the projection, the only
    option at G=14).  Returns logits [B,T,GH,GW]."""
    if net is None:
        net = {}
    if emb is None:
        emb = project_c3d(c3d, p)
    net["c3d_embedded"] = emb
    B, T, G = emb.shape[0], emb.shape[1], emb.shape[2]
    lstm = model == "gaze_lstm"
    ch = (p["ConvLSTM_Wci"] if lstm else p["GRU_Conv_Uz"]).shape[-1]
    state = emb.new_zeros(B, G, G, 2 * ch if lstm else ch)  # zero_state, gaze_grcn.py:132-146
    outs = []
    for i in range(T):  # gaze_grcn.py:269-288
        x = emb[:, i]
        if lstm:
            out, state = lstm_rcn_cell(x, state, p)
        else:
            out, state = gru_rcn_cell(x, state, p)
        outs.append(out)
    net["rcn_outputs"] = outs
    maps = []
    for i in range(T):
        if model == "gaze_grcn77" or "Upsampling/weight1" not in p:
            # gaze_grcn77.py:206-211: per-position FC Ch->1
            o = outs[i].reshape(-1, ch) @ p["out_W"] + p["out_b"]
            maps.append(o.reshape(B, G, G))
        else:
            maps.append(unpool_head(outs[i], p, i, use_bn=(model == "gaze_grcn")))
    logits = torch.stack(maps).permute(1, 0, 2, 3)  # gaze_grcn.py:371-372
    net["predicted_gazemaps"] = logits
    return logits


# ----------------------------------------------------------------------------
# Dense-GRU variant — models/gaze_rnn.py:211-361.  PARITY UNPINNED: the cell arithmetic lives in
# tf.nn.rnn_cell.GRUCell (un-vendored, version unpinned); restated here from the published TF >= 1.2 cell:
#   [r, u] = sigmoid([x, h] . W_gates + b_gates)   (b_gates initialised to 1)
#   c      = tanh([x, r*h] . W_cand + b_cand)      (b_cand initialised to 0)
#   h'     = u*h + (1-u)*c
# The ShallowNet frame-saliency branch (:252-276) is built by the reference but never reaches the prediction
# (its concat entry is commented out, :333), so it is not part of this path.
# ----------------------------------------------------------------------------
def dense_gru_cell(x, h, p):
    H = h.shape[1]
    gates = torch.sigmoid(torch.cat([x, h], 1) @ p["RNN/gru_cell/gates/kernel"] + p["RNN/gru_cell/gates/bias"])
    r, u = gates[:, :H], gates[:, H:]
    c = torch.tanh(torch.cat([x, r * h], 1) @ p["RNN/gru_cell/candidate/kernel"] + p["RNN/gru_cell/candidate/bias"])
    return u * h + (1.0 - u) * c


def dense_gru_network(p, c3d, GH=None, GW=None):
    """gaze_rnn.py:283-359: 1x1 projection 1024 -> dim_cnn_proj (32) per position, the [7,7,32] map flattened (y, x, c
    order) as the input of one dense GRU of 7*7*32 + 49 = 1617 units, proj_out [1617, GH*GW] per step."""
    B, T = c3d.shape[:2]
    if GH is None:
        GH = GW = int(round(math.sqrt(p["proj_out_W"].shape[1])))
    emb = project_c3d(c3d, p)                                   # [B,T,7,7,32]
    H = p["RNN/gru_cell/candidate/bias"].shape[0]
    h = emb.new_zeros(B, H)
    maps = []
    for i in range(T):
        h = dense_gru_cell(emb[:, i].reshape(B, -1), h, p)
        maps.append((h @ p["proj_out_W"] + p["proj_out_b"]).reshape(B, GH, GW))
    return torch.stack(maps).permute(1, 0, 2, 3)


def init_params_dense(dim_proj=32, grid=7, dim_feature=1024, GH=49, GW=49, seed=0, kernel_scale=1.0):
    """Variables of the dense variant in creation order (gaze_rnn.py:294-321): proj_c3d_W/b U(-0.1,0.1); the GRUCell's
    gates / candidate kernels (the reference asks for an "orthogonal" initializer by name, :315) and biases (1 / 0);
    proj_out_W U(-0.1,0.1), proj_out_b zeros."""
    rng = np.random.RandomState(seed)
    I = grid * grid * dim_proj
    H = I + grid * grid                                         # gaze_rnn.py:243-246
    p = OrderedDict()
    p["proj_c3d_W"] = rng.uniform(-0.1, 0.1, size=(dim_feature, dim_proj))
    p["proj_c3d_b"] = rng.uniform(-0.1, 0.1, size=(dim_proj,))

    def orth(rows, cols):
        q, r = np.linalg.qr(rng.standard_normal((max(rows, cols), min(rows, cols))))
        q = q * np.sign(np.diag(r))
        return (q if rows >= cols else q.T) * kernel_scale

    p["RNN/gru_cell/gates/kernel"] = orth(I + H, 2 * H)
    p["RNN/gru_cell/gates/bias"] = np.ones(2 * H)
    p["RNN/gru_cell/candidate/kernel"] = orth(I + H, H)
    p["RNN/gru_cell/candidate/bias"] = np.zeros(H)
    p["proj_out_W"] = rng.uniform(-0.1, 0.1, size=(H, GH * GW))
    p["proj_out_b"] = np.zeros(GH * GW)
    return p


# ----------------------------------------------------------------------------
# Cascade (BASELINE config 5) — models/gaze_grcn_cascade.py:189-452.
#
# As written the file cannot build its graph: the top cell is declared with dim_feature = 64 + 1 (:346-347) and fed
# the 64-channel up-sampled map (:379; the concat with the ShallowNet frame saliency that would supply channel 65 is
# commented out, :366-374) - tf.nn.conv2d rejects the channel mismatch.  Two readings, both restated here:
#   (A) default: the top cell takes the 64 channels it is fed (dim_feature = 64);
#   (B) frm_sal given: the commented-out concat is live, the cell sees [up-sampled 64 | frame saliency 1] = 65
#       channels (the ShallowNet that produces frm_sal, models/saliency_shallownet.py, is outside the path: its map
#       is an input here).
# Everything else is literal: bottom ConvGRU 256 units without BN (:286-312), ONE 11x11 / stride-7 SAME transposed
# conv 7 -> 49, 256 -> 64 (:314-336), 5x5 ConvGRU with 3 units on 49x49 (:343-379), per step two maxout FCs
# 7203 -> 4802 -relu-> max of halves 2401 -> 4802 -relu-> max of halves 2401 (:383-420; dropout on fc1 only, identity
# at keep_prob 1), prediction = that vector as [49,49] with NO softmax (:422-424), loss = sum_t l2_loss / (B*T)
# (:427-441).
# ----------------------------------------------------------------------------
def conv2d_transpose_same_strided(x, w, stride):
    """tf.nn.conv2d_transpose(..., strides=[1,s,s,1], padding='SAME') with output = input * s: the gradient of a SAME
    strided conv [TF-lib]: pad_total = k - s (k >= s), pad_before = pad_total // 2;
    out[b, i*s + ki - pad_before, j*s + kj - pad_before, co] += x[b,i,j,ci] * w[ki,kj,co,ci].
    Used at models/gaze_grcn_cascade.py:326-331 (k = 11, s = 7: pad 2 | 2)."""
    kh = w.shape[0]
    pad_total = max(kh - stride, 0)
    pb = pad_total // 2
    y = F.conv_transpose2d(x.permute(0, 3, 1, 2), w.permute(3, 2, 0, 1), stride=stride, padding=0)
    H = x.shape[1] * stride
    y = y[:, :, pb:pb + H, pb:pb + H]
    return y.permute(0, 2, 3, 1)


def maxout_fc(x, W, b):
    """fully_connected(x, 4802, activation_fn=None) -> relu -> split(1, 2, .) -> maximum  (gaze_grcn_cascade.py:388-420)."""
    y = torch.relu(x @ W + b)
    half = y.shape[1] // 2
    return torch.maximum(y[:, :half], y[:, half:])


def cascade_network(p, c3d, frm_sal=None, net=None):
    """create_gazeprediction_network of gaze_grcn_cascade.py:189-452 (keep_prob = 1).  c3d [B,T,1024,7,7];
    frm_sal [B,T,49,49] or None (reading A).  Returns the predicted maps [B,T,49,49] (not logits: no softmax)."""
    if net is None:
        net = {}
    emb = project_c3d(c3d, p)
    B, T = emb.shape[:2]
    bot = {k.split("/", 1)[1]: v for k, v in p.items() if k.startswith("RCNBottom/")}
    top = {k.split("/", 1)[1]: v for k, v in p.items() if k.startswith("RCNGaze/")}
    ch = bot["GRU_Conv_Uz"].shape[-1]
    state = emb.new_zeros(B, 7, 7, ch)
    outs = []
    for i in range(T):                                  # :293-312
        out, state = gru_rcn_cell(emb[:, i], state, bot)
        outs.append(out)
    net["rcn_outputs"] = outs
    ups = [conv2d_transpose_same_strided(o, p["Upsampling/weight"], 7) for o in outs]   # :323-331
    net["rcn_upsampled_outputs"] = ups
    units = top["GRU_Conv_Uz"].shape[-1]
    sg = emb.new_zeros(B, 49, 49, units)
    maps, tops = [], []
    for i in range(T):                                  # :358-424
        x = ups[i] if frm_sal is None else torch.cat([ups[i], frm_sal[:, i, :, :, None]], dim=3)
        og, sg = gru_rcn_cell(x, sg, top)
        tops.append(og)
        m1 = maxout_fc(og.reshape(B, -1), p["LastProjection/fc1/weights"], p["LastProjection/fc1/bias"])
        m2 = maxout_fc(m1, p["LastProjection/fc2/weights"], p["LastProjection/fc2/bias"])
        maps.append(m2.reshape(B, 49, 49))
    net["top_outputs"] = tops
    pred = torch.stack(maps).permute(1, 0, 2, 3)        # :449
    net["predicted_gazemaps"] = pred
    return pred


def cascade_loss(pred, gt):
    """gaze_grcn_cascade.py:427-441: sum over t of tf.nn.l2_loss(pred_t - gt_t) (= sum of squares / 2), / (B*T)."""
    B, T = pred.shape[:2]
    return 0.5 * ((pred - gt) ** 2).sum() / float(B * T)


def _xavier_uniform_fc(rng, fan_in, fan_out):
    """xavier_initializer(uniform=True) on a [fan_in, fan_out] matrix [TF-lib]."""
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=(fan_in, fan_out))


def init_params_cascade(hidden=256, top_in=64, units=3, dim_proj=512, dim_feature=1024, seed=0, regime="R",
                        stddev_scale=None, fc_bias_scale=0.0):
    """Variables of the cascade in the reference's creation order (gaze_grcn_cascade.py): proj_c3d_W/b (:268-269),
    RCNBottom cell (:52-81 via :287), Upsampling/weight [11,11,64,256] Xavier (:317-321), RCNGaze cell 5x5 with
    `top_in` input channels (64 = reading A, 65 = reading B) and 3 units (:346-351), then on the first time step
    LastProjection/fc1 weights [7203,4802] Xavier + bias 0, fc2 weights [2401,4802] + bias 0 (:388-411).
    fc_bias_scale > 0 draws the biases from N(0, scale) AFTER all other variables (parity tests: a zero bias would
    not exercise the bias path)."""
    rng = np.random.RandomState(seed)
    sc = REGIMES[regime][0] if stddev_scale is None else stddev_scale
    p = OrderedDict()
    p["proj_c3d_W"] = rng.uniform(-0.1, 0.1, size=(dim_feature, dim_proj))
    p["proj_c3d_b"] = rng.uniform(-0.1, 0.1, size=(dim_proj,))

    def cell(scope, k, cin, n):
        for nm, ci in (("Wz", cin), ("Uz", n), ("Wr", cin), ("Ur", n), ("W", cin), ("U", n)):
            p[scope + "/GRU_Conv_" + nm] = _trunc_normal(rng, (k, k, ci, n), 1e-4 * sc)

    cell("RCNBottom", 3, dim_proj, hidden)
    p["Upsampling/weight"] = _xavier_uniform_conv(rng, (11, 11, 64, hidden))
    cell("RCNGaze", 5, top_in, units)
    n_in = 49 * 49 * units
    p["LastProjection/fc1/weights"] = _xavier_uniform_fc(rng, n_in, 4802)
    p["LastProjection/fc1/bias"] = np.zeros(4802)
    p["LastProjection/fc2/weights"] = _xavier_uniform_fc(rng, 2401, 4802)
    p["LastProjection/fc2/bias"] = np.zeros(4802)
    if fc_bias_scale > 0:
        p["LastProjection/fc1/bias"] = fc_bias_scale * rng.standard_normal(4802)
        p["LastProjection/fc2/bias"] = fc_bias_scale * rng.standard_normal(4802)
    return p


def softmax_2d(logits):
    """model_util.py:61-64 applied per frame as gaze_rnn.py:149-159."""
    B, T, H, W = logits.shape
    return torch.softmax(logits.reshape(B, T, H * W), dim=-1).reshape(B, T, H, W)


def loss_fn(logits, gt, loss_type="xentropy"):
    """create_loss_and_summary — gaze_rnn.py:363-408.
    xentropy: sum_t sum_b -sum_j gt*log_softmax(logit) / (B*T); l2: sum 0.5*(pred-gt)^2/(B*T).
    'KLD' is exposed as xentropy - H(gt) (the reference branch raises NameError; same gradient)."""
    B, T, H, W = logits.shape
    if loss_type == "l2":
        return 0.5 * ((logits - gt) ** 2).sum() / float(B * T)
    if loss_type in ("xentropy", "KLD"):
        lsm = torch.log_softmax(logits.reshape(B, T, H * W), dim=-1)
        g = gt.reshape(B, T, H * W)
        loss = -(g * lsm).sum() / float(B * T)
        if loss_type == "KLD":
            ent = torch.where(g > 0, g * torch.log(torch.clamp(g, min=1e-300)), torch.zeros_like(g)).sum()
            loss = loss + ent / float(B * T)
        return loss
    raise NotImplementedError(str(loss_type))


# ----------------------------------------------------------------------------
# Parameters
# ----------------------------------------------------------------------------
def _trunc_normal(rng, shape, std):
    """tf.truncated_normal: re-draw samples beyond 2 sigma [TF-lib]."""
    x = rng.standard_normal(size=shape)
    bad = np.abs(x) > 2.0
    while bad.any():
        x[bad] = rng.standard_normal(size=int(bad.sum()))
        bad = np.abs(x) > 2.0
    return x * std


def _xavier_uniform_conv(rng, shape):
    """xavier_initializer_conv2d(uniform=True): fan_in = kh*kw*shape[2], fan_out = kh*kw*shape[3] [TF-lib]."""
    rf = shape[0] * shape[1]
    lim = math.sqrt(6.0 / (rf * shape[2] + rf * shape[3]))
    return rng.uniform(-lim, lim, size=shape)


REGIMES = {"R": (1.0, False), "S": (200.0, True)}


def init_params(model="gaze_grcn", hidden=128, T=35, grid=7, dim_proj=512, dim_feature=1024,
                seed=0, regime="R", head=None, stddev_scale=None, bn_randomize=None):
    """Parameter set with the reference's variable names and shapes, drawn from
    RandomState(seed) IN THE REFERENCE'S VARIABLE-CREATION ORDER (so that
    oracle/gen_golden_model.py, which runs the reference's graph builders over the eager
    shim with the same seed, creates bit-identical values):
    proj_c3d_W, proj_c3d_b (gaze_grcn.py:234-237, U(-0.1,0.1)); cell kernels in the order of
    GRU_RCN_Cell.__init__ (:64-81) / LSTM_RCN_Cell.__init__ (gaze_lstm.py:64-88), truncated
    normal sigma=1e-4; Upsampling/weight1..3 Xavier-uniform (:292-310); out_W, out_b
    U(-0.1,0.1) (:311-314); per-timestep BN (gamma_t, beta_t) (:325).
    regime 'R' = the reference init.  regime 'S' multiplies the truncated-normal sigma by 200
    (gates leave 0.5, errors compound over T) and randomises BN gamma/beta."""
    rng = np.random.RandomState(seed)
    sc, bnr = REGIMES[regime]
    if stddev_scale is not None:
        sc = stddev_scale
    if bn_randomize is not None:
        bnr = bn_randomize
    p = OrderedDict()
    if head is None:
        head = "fc" if model == "gaze_grcn77" else "deconv"
    p["proj_c3d_W"] = rng.uniform(-0.1, 0.1, size=(dim_feature, dim_proj))
    p["proj_c3d_b"] = rng.uniform(-0.1, 0.1, size=(dim_proj,))

    def kern(cin, cout):
        return _trunc_normal(rng, (3, 3, cin, cout), 1e-4 * sc)

    if model == "gaze_lstm":
        for g in "ifco":
            p["ConvLSTM_Wx" + g] = kern(dim_proj, hidden)
            p["ConvLSTM_Wh" + g] = kern(hidden, hidden)
            if g != "c":
                p["ConvLSTM_Wc" + g] = _trunc_normal(rng, (grid, grid, hidden), 1e-4 * sc)
    else:
        for nm, cin in (("Wz", dim_proj), ("Uz", hidden), ("Wr", dim_proj), ("Ur", hidden),
                        ("W", dim_proj), ("U", hidden)):
            p["GRU_Conv_" + nm] = kern(cin, hidden)
    if head == "deconv":
        p["Upsampling/weight1"] = _xavier_uniform_conv(rng, (5, 5, 64, hidden))
        p["Upsampling/weight2"] = _xavier_uniform_conv(rng, (5, 5, 32, 64))
        p["Upsampling/weight3"] = _xavier_uniform_conv(rng, (7, 7, 12, 32))
        p["out_W"] = rng.uniform(-0.1, 0.1, size=(12, 1))
    else:
        p["out_W"] = rng.uniform(-0.1, 0.1, size=(hidden, 1))
    p["out_b"] = rng.uniform(-0.1, 0.1, size=(1,))
    if model == "gaze_grcn" and head == "deconv":
        # tf.layers.batch_normalization called T times with name=None -> T (gamma,beta) pairs
        p["bn_gamma"] = np.ones((T, hidden))
        p["bn_beta"] = np.zeros((T, hidden))
        if bnr:
            for t in range(T):
                p["bn_gamma"][t] = 1.0 + 0.1 * rng.standard_normal(size=hidden)
                p["bn_beta"][t] = 0.1 * rng.standard_normal(size=hidden)
    return p


def to_torch(p, dtype=torch.float64, requires_grad=False):
    out = OrderedDict()
    for k, v in p.items():
        t = torch.tensor(np.asarray(v), dtype=dtype)
        t.requires_grad_(requires_grad)
        out[k] = t
    return out


def forward_loss_grads(model, p_np, gt, c3d=None, emb=None, loss_type="xentropy", dtype=torch.float64):
    """Forward + loss + autograd BPTT. Returns (logits, probs, loss, grads dict) as NumPy arrays.
    ConvLSTM_Whc receives no gradient (dead variable) -> reported as zeros."""
    p = to_torch(p_np, dtype, requires_grad=True)
    tc = None if c3d is None else torch.tensor(c3d, dtype=dtype)
    te = None if emb is None else torch.tensor(emb, dtype=dtype)
    logits = dense_gru_network(p, tc) if model == "gaze_rnn" else recurrent_network(model, p, c3d=tc, emb=te)
    loss = loss_fn(logits, torch.tensor(gt, dtype=dtype), loss_type)
    names = [k for k in p if not (emb is not None and k.startswith("proj_c3d"))]
    grads = torch.autograd.grad(loss, [p[k] for k in names], allow_unused=True)
    g = OrderedDict()
    for k, gr in zip(names, grads):
        g[k] = np.zeros_like(p_np[k]) if gr is None else gr.detach().numpy()
    return (logits.detach().numpy(), softmax_2d(logits).detach().numpy(), float(loss.detach()), g)


# ----------------------------------------------------------------------------
# Optimiser — models/base.py:262-297, models/gaze_rnn.py:435-444
# ----------------------------------------------------------------------------
def learning_rate(step, lr0=1e-4, decay=0.8, decay_steps=500):
    """tf.train.exponential_decay(staircase=True): lr0 * decay^floor(step/decay_steps)."""
    return lr0 * decay ** (step // decay_steps)


def clip_by_global_norm(grads, clip):
    """tf.clip_by_global_norm: g * clip / max(||g||, clip)."""
    norm = math.sqrt(sum(float((np.asarray(g, dtype=np.float64) ** 2).sum()) for g in grads))
    scale = clip / max(norm, clip)
    return [g * scale for g in grads], norm


def adam_step(params, grads, m, v, step_t, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """tf.train.AdamOptimizer [TF-lib]: lr_t = lr*sqrt(1-b2^t)/(1-b1^t);
    m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; theta -= lr_t * m / (sqrt(v) + eps).
    step_t is the 1-based update count."""
    lr_t = lr * math.sqrt(1.0 - beta2 ** step_t) / (1.0 - beta1 ** step_t)
    for i in range(len(params)):
        m[i] = beta1 * m[i] + (1 - beta1) * grads[i]
        v[i] = beta2 * v[i] + (1 - beta2) * grads[i] ** 2
        params[i] = params[i] - lr_t * m[i] / (np.sqrt(v[i]) + eps)
    return params, m, v


def momentum_step(params, grads, accum, lr, momentum=0.9):
    """tf.train.MomentumOptimizer(lr, momentum=0.9) (models/base.py:272-273) [TF-lib ApplyMomentum, use_nesterov=False]:
    accum = momentum * accum + g; theta -= lr * accum."""
    for i in range(len(params)):
        accum[i] = momentum * accum[i] + grads[i]
        params[i] = params[i] - lr * accum[i]
    return params, accum


def rmsprop_step(params, grads, ms, mom, lr, decay=0.9, momentum=0.9, eps=1e-10):
    """tf.train.RMSPropOptimizer(lr, momentum=0.9) (models/base.py:270-271; decay 0.9 and epsilon 1e-10 are TF1's defaults)
    [TF-lib ApplyRMSProp, centered=False]: ms = decay * ms + (1 - decay) * g^2 (the `rms` slot is created with ones);
    mom = momentum * mom + lr * g / sqrt(ms + eps); theta -= mom."""
    for i in range(len(params)):
        ms[i] = decay * ms[i] + (1 - decay) * grads[i] ** 2
        mom[i] = momentum * mom[i] + lr * grads[i] / np.sqrt(ms[i] + eps)
        params[i] = params[i] - mom[i]
    return params, ms, mom


# ----------------------------------------------------------------------------
# Host pre-processing — models/model_util.py:40-58, models/gaze_rnn.py:499-510
# ----------------------------------------------------------------------------
def normalize_probability_map(t):
    t = np.array(t, copy=True)
    if t.ndim == 3:
        for i in range(len(t)):
            t[i, :] /= t[i, :].sum()
    elif t.ndim == 4:
        for i in range(t.shape[0]):
            for j in range(t.shape[1]):
                t[i, j, :] /= t[i, j, :].sum()
    else:
        raise ValueError("Unsupported shape : {}".format(t.shape))
    return t


# ----------------------------------------------------------------------------
# Naive NumPy definitions used only to pin the torch conv semantics above
# ----------------------------------------------------------------------------
def naive_conv2d_same(x, w):
    B, H, W, Ci = x.shape
    kh, kw, _, Co = w.shape
    out = np.zeros((B, H, W, Co))
    for y in range(H):
        for xx in range(W):
            for ky in range(kh):
                for kx in range(kw):
                    sy, sx = y + ky - kh // 2, xx + kx - kw // 2
                    if 0 <= sy < H and 0 <= sx < W:
                        out[:, y, xx, :] += x[:, sy, sx, :] @ w[ky, kx]
    return out


def naive_conv2d_transpose(x, w, stride, padding):
    B, H, W, Ci = x.shape
    kh, kw, Co, _ = w.shape
    OH, OW = (H - 1) * stride + kh, (W - 1) * stride + kw
    out = np.zeros((B, OH, OW, Co))
    for i in range(H):
        for j in range(W):
            for ki in range(kh):
                for kj in range(kw):
                    out[:, i * stride + ki, j * stride + kj, :] += x[:, i, j, :] @ w[ki, kj].T
    if padding == "SAME":
        c = (kh - 1) // 2
        out = out[:, c:c + H, c:c + W, :]
    return out
