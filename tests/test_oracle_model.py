"""The model oracle against goldens minted by RUNNING the reference's graph builders
(oracle/gen_golden_model.py: models/gaze_grcn.py, gaze_grcn77.py, gaze_lstm.py, gaze_rnn.py,
model_util.py imported unmodified, TF ops served by oracle/tf1_eager_shim.py).  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import synthetic as SY

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(model):
    return np.load(os.path.join(GOLD, "model_ref_%s.npz" % model))


def _inputs(g, GH, GW):
    rng = np.random.RandomState(int(g["input_seed"]))
    B, T = int(g["B"]), int(g["T"])
    c3d = SY.make_c3d(rng, B, T).astype(np.float64)
    gt, _ = SY.make_gazemaps(rng, B, T, GH, GW, normalize=True)
    gt = gt.astype(np.float64)
    gt /= gt.sum(axis=(2, 3), keepdims=True)
    return c3d, gt


def _get(p, name):
    if name.startswith("bn_"):
        base, t = name[:-1].split("[")
        return p[base][int(t)]
    return p[name]


@pytest.mark.parametrize("model", ["gaze_grcn", "gaze_grcn77", "gaze_lstm"])
def test_oracle_matches_reference_graph_run(model):
    g = _load(model)
    GH = 7 if model == "gaze_grcn77" else 49
    c3d, gt = _inputs(g, GH, GH)
    p = O.init_params(model, hidden=128, T=int(g["T"]), seed=int(g["seed"]),
                      stddev_scale=float(g["stddev_scale"]), bn_randomize=True)
    names = [str(n) for n in g["var_names"]]
    # the oracle's parameter draw reproduces the reference's variable creation bit-for-bit
    for nm in names:
        a = np.asarray(_get(p, nm), dtype=np.float64).ravel()
        assert np.array_equal(a[g["var/%s/idx" % nm]], g["var/%s/samp" % nm]), nm
        assert abs(a.sum() - g["var/%s/sum" % nm][0]) <= 1e-9 * max(1, abs(a.sum()))
    for lt in ("xentropy", "l2"):
        logits, probs, loss, grads = O.forward_loss_grads(model, p, gt, c3d=c3d, loss_type=lt)
        assert np.allclose(logits, g["logits"], rtol=1e-10, atol=1e-12)
        assert np.allclose(probs, g["probs"], rtol=1e-10, atol=1e-14)
        assert abs(loss - float(g["loss_" + lt])) <= 1e-10 * abs(loss)
        tag = "gx" if lt == "xentropy" else "gl2"
        for nm in names:
            a = np.asarray(_get(grads, nm), dtype=np.float64).ravel()
            ref_sum, ref_l2 = g["%s/%s/sum" % (tag, nm)]
            assert np.allclose(a[g["var/%s/idx" % nm]], g["%s/%s/samp" % (tag, nm)], rtol=1e-8, atol=1e-14), nm
            assert abs(np.sqrt((a * a).sum()) - ref_l2) <= 1e-8 * max(ref_l2, 1e-30), nm
            if bool(g["%s/%s/none" % (tag, nm)]):
                assert nm == "ConvLSTM_Whc" and not a.any()  # dead variable


def test_conv_semantics_against_naive_definitions():
    rng = np.random.RandomState(0)
    x = rng.standard_normal(size=(2, 5, 5, 3))
    w = rng.standard_normal(size=(3, 3, 3, 4))
    a = O.conv2d_same(torch.tensor(x), torch.tensor(w)).numpy()
    assert np.allclose(a, O.naive_conv2d_same(x, w), atol=1e-12)
    for (k, s, pad, co) in ((5, 3, "VALID", 6), (5, 2, "VALID", 2), (7, 1, "SAME", 3)):
        wt = rng.standard_normal(size=(k, k, co, 3))
        a = O.conv2d_transpose(torch.tensor(x), torch.tensor(wt), s, pad).numpy()
        b = O.naive_conv2d_transpose(x, wt, s, pad)
        assert a.shape == b.shape and np.allclose(a, b, atol=1e-12)
    # the reference's asserted head shapes 7 -> 23 -> 49 -> 49 (gaze_grcn.py:329,341,356)
    h = torch.zeros(1, 7, 7, 8, dtype=torch.float64)
    y = O.conv2d_transpose(h, torch.zeros(5, 5, 4, 8, dtype=torch.float64), 3, "VALID")
    assert y.shape[1:3] == (23, 23)
    y = O.conv2d_transpose(y, torch.zeros(5, 5, 2, 4, dtype=torch.float64), 2, "VALID")
    assert y.shape[1:3] == (49, 49)


def test_gru_update_convention_and_zero_state():
    """z=1 keeps the old state; h_0 = 0."""
    p = O.to_torch(O.init_params("gaze_grcn77", hidden=4, dim_proj=3, dim_feature=5, seed=0))
    x = torch.zeros(1, 7, 7, 3, dtype=torch.float64)
    h = torch.ones(1, 7, 7, 4, dtype=torch.float64)
    for k in p:
        if k.startswith("GRU_Conv"):
            p[k] = torch.zeros_like(p[k])
    out, _ = O.gru_rcn_cell(x, h, p)  # z = 0.5, c = 0 -> h' = 0.5 h
    assert torch.allclose(out, 0.5 * h)


def test_optimizer_restatement():
    rng = np.random.RandomState(0)
    g = [rng.standard_normal(size=(4, 5)) * 10, rng.standard_normal(size=(7,)) * 10]
    cl, n = O.clip_by_global_norm(g, 10.0)
    assert abs(np.sqrt(sum((c ** 2).sum() for c in cl)) - 10.0) < 1e-12 and n > 10
    cl2, _ = O.clip_by_global_norm([x * 1e-3 for x in g], 10.0)
    assert np.array_equal(cl2[0], g[0] * 1e-3)
    assert O.learning_rate(499) == 1e-4 and abs(O.learning_rate(500) - 0.8e-4) < 1e-20
    assert abs(O.learning_rate(1234) - 1e-4 * 0.8 ** 2) < 1e-20
    # TF-Adam (epsilon outside the bias correction) vs the textbook form differ
    p0 = [np.ones(3)]
    p1, m, v = O.adam_step([p0[0].copy()], [np.full(3, 1e-9)], [np.zeros(3)], [np.zeros(3)], 1, 1e-3)
    lr_t = 1e-3 * np.sqrt(1 - 0.999) / (1 - 0.9)
    assert np.allclose(p1[0], 1 - lr_t * (0.1e-9) / (np.sqrt(0.001 * 1e-18) + 1e-8))


def test_normalize_probability_map():
    rng = np.random.RandomState(0)
    t = rng.uniform(size=(2, 3, 4, 4)).astype(np.float32)
    n = O.normalize_probability_map(t)
    assert np.allclose(n.sum(axis=(2, 3)), 1, atol=1e-6) and n.dtype == np.float32


def test_cascade_oracle_matches_reference_graph_run():
    """oracle.cascade_network against the reference's gaze_grcn_cascade.py create_gazeprediction_network run over the
    shim (oracle/gen_golden_model.py run_cascade: reading A of the 65-vs-64 channel bug): variable draw bit-for-bit,
    bottom states, up-sampled maps, predicted maps and the l2 loss."""
    g = _load("gaze_grcn_cascade")
    B, T = int(g["B"]), int(g["T"])
    rng = np.random.RandomState(int(g["input_seed"]))
    c3d = SY.make_c3d(rng, B, T).astype(np.float64)
    gt, _ = SY.make_gazemaps(rng, B, T, 49, 49, normalize=False)
    p = O.init_params_cascade(seed=int(g["seed"]), stddev_scale=float(g["stddev_scale"]))
    names = [str(n) for n in g["var_names"]]
    assert names == list(p.keys())                       # same variables, same creation order
    for nm in names:
        a = np.asarray(p[nm], dtype=np.float64)
        assert list(a.shape) == list(g["var/%s/shape" % nm]), nm
        a = a.ravel()
        assert np.array_equal(a[g["var/%s/idx" % nm]], g["var/%s/samp" % nm]), nm
        assert abs(a.sum() - g["var/%s/sum" % nm][0]) <= 1e-9 * max(1, abs(a.sum()))
    net = {}
    with torch.no_grad():
        pred = O.cascade_network(O.to_torch(p), torch.tensor(c3d), net=net)
        loss = float(O.cascade_loss(pred, torch.tensor(gt.astype(np.float64))))
    assert np.allclose(net["rcn_outputs"][-1].numpy(), g["rcn_last"], rtol=1e-10, atol=1e-14)
    up = net["rcn_upsampled_outputs"][-1].numpy().ravel()
    assert abs(up.sum() - g["up_fp"][0]) <= 1e-9 * max(1.0, abs(up.sum()))
    assert abs(np.sqrt((up * up).sum()) - g["up_fp"][1]) <= 1e-10 * g["up_fp"][1]
    assert np.allclose(pred.numpy(), g["pred"], rtol=1e-10, atol=1e-13)
    assert pred.numpy().max() > 1e-3                      # the maxout chain is alive at this scale
    assert abs(loss - float(g["loss_l2"])) <= 1e-10 * abs(loss)


def test_strided_same_transposed_conv_against_naive_definition():
    """tf.nn.conv2d_transpose(k=11, stride 7, SAME) 7 -> 49 (gaze_grcn_cascade.py:326-331)."""
    rng = np.random.RandomState(0)
    x = rng.standard_normal((1, 7, 7, 3))
    w = rng.standard_normal((11, 11, 2, 3))
    y = O.conv2d_transpose_same_strided(torch.tensor(x), torch.tensor(w), 7).numpy()
    out = np.zeros((1, 49, 49, 2))
    for i in range(7):
        for j in range(7):
            for ki in range(11):
                for kj in range(11):
                    oy, ox = i * 7 + ki - 2, j * 7 + kj - 2
                    if 0 <= oy < 49 and 0 <= ox < 49:
                        out[:, oy, ox, :] += x[:, i, j, :] @ w[ki, kj].T
    assert np.allclose(y, out, atol=1e-12)
