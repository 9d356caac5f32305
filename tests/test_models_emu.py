"""The reference-facing model classes (recurrent_gaze_prediction_b200/models/*) over the NumPy emulation of the
C-ABI: constructor / attribute / method contract of gaze_rnn.py, cells vs the oracle's cell maths, train steps vs the
oracle's clip + TF-Adam, checkpoint round trip, generate/evaluate dict contract.  CPU only."""
import os

import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.models import gaze_grcn, gaze_grcn77, gaze_lstm, gaze_rnn
from recurrent_gaze_prediction_b200.models.model_util import init_parameters
from emu_lib import EmuLib


def small_config(T=3, B=4, Ch=16):
    c = gaze_rnn.GRUModelConfig()
    c.n_lstm_steps, c.batch_size, c.rnn_state_size = T, B, Ch
    c.dim_feature, c.dim_cnn_proj = 64, 32
    c.dtype, c.device = 'fp32', 'cpu'
    c.initial_learning_rate = 1e-2
    c.steps_per_logprint = 1000
    return c


def data(T=3, n=12, **kw):
    return SY.read_synthetic_data_sets(n, 12, 12, T=T, dim_feature=64, **kw)


def test_constructor_contract_and_single_step_matches_oracle():
    c = small_config()
    ds = data()
    m = gaze_grcn.GazePredictionGRCN({'lib': EmuLib()}, ds, c)
    for attr in ("batch_size", "n_lstm_steps", "predicted_gazemaps", "predicted_gazemaps_logit", "loss", "c3d_input",
                 "frame_images", "gt_gazemap", "dropout_keep_prob", "train_op", "global_step", "learning_rate", "session"):
        assert hasattr(m, attr), attr
    assert tuple(m.c3d_input.shape) == (4, 3, 64, 7, 7) and tuple(m.gt_gazemap.shape) == (4, 3, 49, 49)
    assert m.current_step == 0 and abs(m.current_learning_rate - 1e-2) < 1e-12
    # oracle replay of the same step: same parameters (same seed, same creation order), same batch, same flip draw
    p = init_parameters("gaze_grcn", hidden=16, T=3, dim_proj=32, dim_feature=64, seed=0)
    ref_ds = data()
    imgs, maps, fix, c3d, _, _ = ref_ds.train.next_batch(4)
    c3d = c3d.reshape(4, 3, 64, 7, 7)
    maps = O.normalize_probability_map(maps)
    np.random.seed(5)
    idx = np.random.choice(4, 2, replace=False)
    maps[idx] = maps[idx][:, :, :, ::-1]
    c3d[idx] = c3d[idx][:, :, :, :, ::-1]
    _, _, ref_loss, grads = O.forward_loss_grads("gaze_grcn", p, maps.astype(np.float64), c3d=c3d.astype(np.float64))
    np.random.seed(5)
    step = m.single_step(train_mode=True)
    assert step == 1 and m.current_step == 1
    assert abs(m.last_loss - ref_loss) <= 1e-5 * abs(ref_loss)
    names = list(p)
    gl, _ = O.clip_by_global_norm([grads[k] for k in names], 10.0)
    newp, _, _ = O.adam_step([np.asarray(p[k], np.float64) for k in names], gl, [np.zeros_like(p[k]) for k in names],
                             [np.zeros_like(p[k]) for k in names], 1, 1e-2)
    got = m.engine.op_to_tf(m.engine.params.numpy())
    for k, r in zip(names, newp):
        if k == "out_b":
            continue    # its true gradient is sum(softmax - gt) = 0: Adam normalises rounding noise to +-lr
        assert np.allclose(got[k].reshape(r.shape), r, rtol=2e-4, atol=1e-6), k
    # validation step does not train
    before = m.engine.params.clone()
    assert m.single_step(train_mode=False) == 1
    assert torch.equal(before, m.engine.params)


def test_staircase_learning_rate_and_decay():
    c = small_config()
    m = gaze_grcn.GazePredictionGRCN({'lib': EmuLib()}, data(), c)
    assert abs(m._learning_rate_at(499) - 1e-2) < 1e-15
    assert abs(m._learning_rate_at(500) - 1e-2 * 0.8) < 1e-15
    assert abs(m._learning_rate_at(1234) - 1e-2 * 0.8 ** 2) < 1e-15
    assert abs(m.decay_learning_rate(0.5) - 5e-3) < 1e-15


def test_checkpoint_roundtrip(tmp_path):
    c = small_config()
    c.train_dir = str(tmp_path)
    lib = EmuLib()
    m = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c)
    np.random.seed(0)
    m.single_step()
    m.single_step()
    path = m.save_model_checkpoint(c.train_dir)
    assert os.path.isfile(path) and os.path.isfile(os.path.join(c.train_dir, "model", "checkpoint"))
    with np.load(path) as z:
        assert "GRU_Conv_Wz" in z.files and z["GRU_Conv_Wz"].shape == (3, 3, 32, 16)      # reference names + TF layouts
        assert "Upsampling/weight1" in z.files and int(z["__global_step__"]) == 2
    c2 = small_config()
    c2.train_dir, c2.seed = str(tmp_path), 123
    m2 = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c2)
    assert not torch.equal(m.engine.params, m2.engine.params)
    assert m2.reload_checkpoint()
    assert torch.allclose(m.engine.params, m2.engine.params) and m2.current_step == 2
    assert torch.equal(m.engine.adam_v, m2.engine.adam_v)
    c3 = small_config()
    c3.train_dir = str(tmp_path / "nothing_here")
    assert not gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c3).reload_checkpoint()


def test_generate_and_evaluate_contract():
    c = small_config(T=3, B=4)
    ds = data(n=12)
    m = gaze_grcn.GazePredictionGRCN({'lib': EmuLib()}, ds, c)
    ret = m.generate(ds.valid, max_instances=8)
    assert sorted(ret) == ['c3d_list', 'clipname_list', 'fixationmap_list', 'gt_gazemap_list', 'images_list',
                           'pred_gazemap_list']
    assert ret['pred_gazemap_list'].shape == (8 * 3, 49, 49) and ret['gt_gazemap_list'].shape == (24, 49, 49)
    assert np.allclose(ret['pred_gazemap_list'].sum(axis=(1, 2)), 1.0, atol=1e-4)          # probabilities (gaze_rnn.py:156)
    assert len(ret['fixationmap_list']) == 24 and len(ret['images_list']) == 24 and len(ret['clipname_list']) == 8
    np.random.seed(1)
    scores = m.evaluate(**ret)
    assert sorted(scores) == sorted(['sim', 'cc', 'AUC_shuffled', 'AUC_Borji'])
    assert 0 < scores['sim'] <= 1 and -1 <= scores['cc'] <= 1 and 0 <= scores['AUC_Borji'] <= 1
    assert m.evaluation_history[-1][1] == scores
    with pytest.raises(AssertionError):
        m.evaluate(ret['pred_gazemap_list'][:3], ret['gt_gazemap_list'], ret['fixationmap_list'], ret['images_list'])


def test_loss_builder_and_errors():
    lib = EmuLib()
    rng = np.random.RandomState(0)
    logits = torch.from_numpy(rng.standard_normal((2, 3, 7, 7)).astype(np.float32))
    gt = torch.from_numpy(O.normalize_probability_map(rng.uniform(size=(2, 3, 7, 7))).astype(np.float32))
    for lt in ("xentropy", "l2", "KLD"):
        got = float(gaze_rnn.GazePredictionGRU.create_loss_and_summary(logits, gt, lt, lib=lib)[0])
        want = float(O.loss_fn(logits.double(), gt.double(), lt))
        assert abs(got - want) <= 1e-5 * abs(want) + 1e-7, lt
    with pytest.raises(NotImplementedError):
        gaze_rnn.GazePredictionGRU.create_loss_and_summary(logits, gt, "hinge", lib=lib)
    c = small_config()
    c.loss_type = "hinge"
    with pytest.raises(NotImplementedError):
        gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c)
    c = small_config(B=4)
    with pytest.raises(ValueError):
        gaze_grcn.GazePredictionGRCN({'lib': lib, 'world_size': 3}, data(), c)


def test_cells_match_oracle_cell_maths():
    lib = EmuLib()
    rng = np.random.RandomState(3)
    B, G, Cin, Ch = 2, 7, 32, 16
    x = torch.from_numpy(rng.standard_normal((B, G, G, Cin)).astype(np.float32))
    # GRU cell
    cell = gaze_grcn.GRU_RCN_Cell(Ch, Cin, device='cpu', lib=lib)
    assert (cell.input_size, cell.output_size, cell.state_size) == (Cin, Ch, Ch)
    big = {k: v * 2000 for k, v in cell.variables.items()}          # leave the sigma=1e-4 regime so gates move
    cell.load_variables(big)
    h0 = torch.from_numpy(rng.standard_normal((B, G, G, Ch)).astype(np.float32))
    out, h1 = cell(x, h0)
    pt = {k: torch.tensor(v, dtype=torch.float64) for k, v in big.items()}
    _, ref = O.gru_rcn_cell(x.double(), h0.double(), pt)
    assert torch.allclose(h1.double(), ref, atol=2e-5) and torch.equal(out, h1)
    out2, h2 = cell(x, cell.zero_state(B))
    _, ref2 = O.gru_rcn_cell(x.double(), torch.zeros(B, G, G, Ch, dtype=torch.float64), pt)
    assert torch.allclose(h2.double(), ref2, atol=2e-5)
    with pytest.raises(AssertionError):
        cell(torch.zeros(B, 6, 7, Cin), h0)
    # LSTM cell (reference arithmetic as written)
    lc = gaze_lstm.LSTM_RCN_Cell(Ch, Cin, device='cpu', lib=lib)
    assert lc.state_size == 2 * Ch
    bigl = {k: v * 2000 for k, v in lc.variables.items()}
    lc.load_variables(bigl)
    s0 = torch.from_numpy(rng.standard_normal((B, G, G, 2 * Ch)).astype(np.float32))
    h, s1 = lc(x, s0)
    ptl = {k: torch.tensor(v, dtype=torch.float64) for k, v in bigl.items()}
    refh, refs = O.lstm_rcn_cell(x.double(), s0.double(), ptl)
    assert torch.allclose(s1.double(), refs, atol=2e-5) and torch.allclose(h.double(), refh, atol=2e-5)


@pytest.mark.parametrize("mod,cls,model,GH", [(gaze_grcn, "GazePredictionGRCN", "gaze_grcn", 49),
                                              (gaze_grcn77, "GazePredictionGRCN", "gaze_grcn77", 7),
                                              (gaze_lstm, "GazePredictionLSTM", "gaze_lstm", 49)])
def test_network_builders_match_oracle(mod, cls, model, GH):
    rng = np.random.RandomState(1)
    B, T = 2, 3
    c3d = SY.make_c3d(rng, B, T, 64, 7)
    p = O.init_params(model, hidden=16, T=T, dim_proj=32, dim_feature=64, seed=4, regime="S")
    net = {}
    logits = getattr(mod, cls).create_gazeprediction_network(None, torch.from_numpy(c3d), net=net, params=p,
                                                             rnn_state_size=16, lib=EmuLib())
    ref, _, _, _ = O.forward_loss_grads(model, p, np.full((B, T, GH, GH), 1.0 / GH ** 2), c3d=c3d.astype(np.float64))
    assert tuple(logits.shape) == (B, T, GH, GH)
    assert np.abs(logits.numpy() - ref).max() <= 2e-5 * np.abs(ref).max()
    assert tuple(net['c3d_embedded'].shape) == (B, T, 7, 7, 32) and len(net['rcn_outputs']) == T
    onet = {}
    O.recurrent_network(model, O.to_torch(p), c3d=torch.from_numpy(c3d).double(), net=onet)
    assert torch.allclose(net['c3d_embedded'].double(), onet['c3d_embedded'], atol=1e-4)
    assert torch.allclose(net['rcn_outputs'][-1].double(), onet['rcn_outputs'][-1], atol=1e-4)


@pytest.mark.parametrize("Model", [gaze_grcn.GazePredictionGRCN, gaze_lstm.GazePredictionLSTM])
def test_chunked_long_clip_equals_one_pass(Model, tmp_path):
    """extract_map.py:65,148-238: long clips run `chunk_steps` frames at a time with the state (and the per-time-step
    BN parameters' time offset) carried between chunks; the gaze maps equal the one-pass model's."""
    from recurrent_gaze_prediction_b200.models import extract_map as XM
    lib = EmuLib()
    T, B = 6, 2
    c1, c2 = small_config(T=T, B=B), small_config(T=T, B=B)
    c1.trainable = c2.trainable = False
    c2.chunk_steps = 2
    one = Model({'lib': lib}, data(T=T), c1)
    chk = Model({'lib': lib}, data(T=T), c2)
    assert chk.engine.T == 2 and chk.engine.Tp == T and one.engine.T == T
    if one.engine.use_bn:          # distinct BN parameters per time step, so a wrong time offset would show
        g = one.engine.view("bn_gamma")
        g.copy_(torch.linspace(0.5, 1.5, g.numel()).view_as(g))
        one.engine.view("bn_beta").copy_(torch.linspace(-0.2, 0.2, g.numel()).view_as(g))
    chk.engine.params.copy_(one.engine.params)
    chk.engine.refresh_operands()
    one.engine.refresh_operands()
    rng = np.random.RandomState(3)
    clips = [("a", SY.make_c3d(rng, 1, 4, dim_feature=64)[0]), ("b", SY.make_c3d(rng, 1, 6, dim_feature=64)[0]),
             ("c", SY.make_c3d(rng, 1, 9, dim_feature=64)[0])]           # short (padded), exact, long (cut)
    r1 = XM.extract(one, clips)
    r2 = XM.extract(chk, clips, out_dir=str(tmp_path))
    for name, n in (("a", 4), ("b", 6), ("c", 6)):
        g49, g7 = r1[name]
        assert g49.shape == (n, 49, 49) and g7.shape == (n, 7, 7)
        assert np.allclose(g49.sum(axis=(1, 2)), 1.0, atol=1e-5)                     # probabilities
        assert np.allclose(r2[name][0], g49, rtol=1e-6, atol=1e-9), name
        assert np.allclose(g7, g49.reshape(n, 7, 7, 7, 7).mean(axis=(2, 4)))
        assert np.array_equal(np.load(os.path.join(str(tmp_path), name, name + ".gazemap.49.npy")), r2[name][0])
        assert np.load(os.path.join(str(tmp_path), name, name + ".gazemap.npy")).shape == (n, 7, 7)
    # a second batch through the same chunked model starts from h_0 = 0 again
    again = XM.extract(chk, clips[:1])
    assert np.array_equal(again["a"][0], r2["a"][0])
    with pytest.raises(ValueError):
        bad = small_config(T=T, B=B)
        bad.chunk_steps = 2                    # trainable model
        Model({'lib': lib}, data(T=T), bad)


def test_dense_gru_variant_matches_oracle():
    """SURVEY row a24 (gaze_rnn.py:211-361): the class named GazePredictionGRU is the dense GRU variant; forward, loss,
    checkpoint layout and generate() against the oracle's restatement of tf.nn.rnn_cell.GRUCell (parity unpinned at
    that library boundary)."""
    from recurrent_gaze_prediction_b200.dense_engine import DenseGruEngine
    rng = np.random.RandomState(0)
    B, T = 3, 4
    p = O.init_params_dense(dim_proj=4, dim_feature=16, GH=7, GW=7, seed=2, kernel_scale=1.5)
    p["RNN/gru_cell/gates/bias"] = 1.0 + 0.3 * rng.standard_normal(p["RNN/gru_cell/gates/bias"].shape)
    p["RNN/gru_cell/candidate/bias"] = 0.3 * rng.standard_normal(p["RNN/gru_cell/candidate/bias"].shape)
    p["proj_out_b"] = 0.1 * rng.standard_normal(p["proj_out_b"].shape)
    e = DenseGruEngine(B=B, T=T, Cp=4, Cfeat=16, GH=7, dtype=0, device="cpu", lib=EmuLib())
    assert (e.I, e.H, e.Ip, e.Hp, e.Np) == (196, 245, 256, 256, 128)
    e.load_tf_params(p)
    back = e.op_to_tf(e.params.numpy())
    for k in p:
        assert np.allclose(back[k], np.asarray(p[k], np.float32)), k
    c3d = SY.make_c3d(rng, B, T, dim_feature=16)
    gt = O.normalize_probability_map(SY.make_gazemaps(rng, B, T, 7, 7)[0]).astype(np.float32)
    logits = e.forward(c3d=torch.from_numpy(c3d), gt=torch.from_numpy(gt)).numpy()
    ref, ref_probs, ref_loss, _ = O.forward_loss_grads("gaze_rnn", p, gt.astype(np.float64), c3d=c3d.astype(np.float64))
    assert np.abs(logits - ref).max() <= 1e-4 * max(1.0, np.abs(ref).max())
    assert np.abs(e.probs.numpy() - ref_probs).max() <= 1e-5
    assert abs(float(e.loss[0]) - ref_loss) <= 1e-5 * abs(ref_loss)
    assert np.all(e.Hf.numpy()[:, e.H:] == 0.0)                  # padded state columns stay exactly zero
    # real sizes of the reference: 7*7*32 + 49 = 1617 units, padded 1600 / 1664 / 2432
    big = DenseGruEngine.__new__(DenseGruEngine)
    big.P, big.Cp = 49, 32
    assert (49 * 32 + 49, (49 * 32 + 63) // 64 * 64, (1617 + 127) // 128 * 128, (2401 + 127) // 128 * 128) == (1617, 1600, 1664, 2432)


@pytest.mark.parametrize("keep_prob", [1.0, 0.5])
def test_dense_gru_bptt_matches_oracle_autograd(keep_prob):
    """BPTT of the dense variant (dense_engine.backward) against torch autograd on the oracle's restatement; with dropout
    the engine's own counter-based mask (read back from the packed input rows) is replayed on the oracle."""
    from recurrent_gaze_prediction_b200.dense_engine import DenseGruEngine
    rng = np.random.RandomState(1)
    B, T = 3, 4
    p = O.init_params_dense(dim_proj=4, dim_feature=16, GH=7, GW=7, seed=2, kernel_scale=1.5)
    p["RNN/gru_cell/gates/bias"] = 1.0 + 0.3 * rng.standard_normal(p["RNN/gru_cell/gates/bias"].shape)
    p["RNN/gru_cell/candidate/bias"] = 0.3 * rng.standard_normal(p["RNN/gru_cell/candidate/bias"].shape)
    e = DenseGruEngine(B=B, T=T, Cp=4, Cfeat=16, GH=7, dtype=0, device="cpu", lib=EmuLib(), train=True)
    e.load_tf_params(p)
    e.keep_prob, e.dropout_seed = keep_prob, 12345
    c3d = SY.make_c3d(rng, B, T, dim_feature=16)
    gt = O.normalize_probability_map(SY.make_gazemaps(rng, B, T, 7, 7)[0]).astype(np.float32)
    e.forward(c3d=torch.from_numpy(c3d), gt=torch.from_numpy(gt))
    e.backward()
    got = e.op_to_tf(e.grads.numpy())
    # oracle with the same mask on the embedded features
    pt = O.to_torch(p, requires_grad=True)
    emb = O.project_c3d(torch.tensor(c3d, dtype=torch.float64), pt)               # [B,T,7,7,4]
    X = e.X.numpy()[:T * B, :e.I].reshape(T, B, 7, 7, 4).transpose(1, 0, 2, 3, 4)
    if keep_prob < 1.0:
        mask = torch.tensor((X != 0).astype(np.float64)) / keep_prob
        frac = float((X != 0).mean())
        assert 0.4 < frac < 0.6
        emb = emb * mask
    H = p["RNN/gru_cell/candidate/bias"].shape[0]
    h = emb.new_zeros(B, H)
    maps = []
    for i in range(T):
        h = O.dense_gru_cell(emb[:, i].reshape(B, -1), h, pt)
        maps.append((h @ pt["proj_out_W"] + pt["proj_out_b"]).reshape(B, 7, 7))
    logits = torch.stack(maps).permute(1, 0, 2, 3)
    loss = O.loss_fn(logits, torch.tensor(gt, dtype=torch.float64))
    assert abs(float(e.loss[0]) - float(loss)) <= 1e-5 * abs(float(loss))
    grads = torch.autograd.grad(loss, [pt[k] for k in p])
    for k, gr in zip(p, grads):
        r = gr.numpy()
        assert np.abs(got[k] - r).max() <= 2e-4 * max(np.abs(r).max(), 1e-12), k
    # nothing leaks into the zero padding of the op layouts
    assert np.all(e.grads.numpy()[e.pad_mask() == 0] == 0.0)


def test_dense_gru_model_class_contract():
    c = small_config(T=3, B=2)
    c.dim_cnn_proj_dense = 4
    c.trainable = False
    ds = data(T=3)
    m = gaze_rnn.GazePredictionGRU({'lib': EmuLib()}, ds, c)
    assert type(m.engine).__name__ == "DenseGruEngine" and m.train_op is None
    assert tuple(m.predicted_gazemaps.shape) == (2, 3, 49, 49)
    assert m.single_step(train_mode=False) == 0 and np.isfinite(m.last_loss)
    with pytest.raises(RuntimeError):
        m.single_step(train_mode=True)
    ret = m.generate(ds.valid, max_instances=2)
    assert ret['pred_gazemap_list'].shape == (6, 49, 49)
    assert np.allclose(ret['pred_gazemap_list'].sum(axis=(1, 2)), 1.0, atol=1e-5)
    # trainable (the reference trains this class: gaze_rnn.py:447-478): steps run, dropout live, loss finite and moving
    c.trainable = True
    c.initial_learning_rate = 1e-2
    mt = gaze_rnn.GazePredictionGRU({'lib': EmuLib()}, ds, c)
    assert mt.train_op is not None
    p0 = mt.engine.params.clone()
    for k in range(2):
        assert mt.single_step(train_mode=True) == k + 1
    assert mt.engine.keep_prob == 0.5 and np.isfinite(mt.last_loss)
    moved = (mt.engine.params - p0).abs().numpy()
    mask = mt.engine.pad_mask()
    assert moved[mask > 0].max() > 1e-4 and moved[mask == 0].max() == 0.0        # the zero padding never moves
    mt.single_step(train_mode=False)
    assert mt.engine.keep_prob == 1.0


def test_optimisation_method_and_device_side_schedule(tmp_path):
    """base.py:268-275: an unknown optimisation method is a ValueError; 'rmsprop' (RMSProp, momentum 0.9) and 'sgd'
    (Momentum 0.9) take TF1's update rules on the globally clipped gradient (oracle: rmsprop_step / momentum_step, the rms
    slot starting at one).  The staircase learning rate and Adam's bias correction are evaluated by rgp_adam_schedule from
    the engine's own step counter, with the config's decay interval."""
    lib = EmuLib()
    c = small_config()
    c.optimization_method = "adagrad"
    with pytest.raises(ValueError):
        gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c)
    for method in ("rmsprop", "sgd"):
        c = small_config()
        c.optimization_method = method
        m = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c)
        np.random.seed(1)
        n = m.engine.nparams
        ms = np.ones(n) if method == "rmsprop" else None
        acc = np.zeros(n)
        for k in range(3):
            p0 = m.engine.params.numpy().astype(np.float64).copy()
            m.single_step()
            g = m.engine.grads.numpy().astype(np.float64)
            (gc,), _ = O.clip_by_global_norm([g], m.max_grad_norm)
            lr = float(m.engine.sched[1])
            if method == "rmsprop":
                (want,), (ms,), (acc,) = O.rmsprop_step([p0], [gc], [ms], [acc], lr)
            else:
                (want,), (acc,) = O.momentum_step([p0], [gc], [acc], lr)
            got = m.engine.params.numpy().astype(np.float64)
            assert np.abs(got - p0).max() > 0
            assert np.abs(got - want).max() <= 2e-6 * max(np.abs(want - p0).max(), 1e-30) + 1e-7 * np.abs(want).max(), (method, k)
        assert int(m.engine.step_dev[0]) == 3 == m.current_step
    c = small_config()
    c.learning_rate_decay_steps, c.learning_rate_decay, c.train_dir = 2, 0.5, str(tmp_path)
    m = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c)
    np.random.seed(0)
    lrs = []
    for k in range(5):
        m.single_step()
        lrs.append(float(m.engine.sched[1]))
        assert int(m.engine.step_dev[0]) == k + 1 == m.current_step
        t = k + 1
        want = lrs[-1] * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        assert abs(float(m.engine.sched[0]) - want) <= 2e-7 * want
    assert np.allclose(lrs, [1e-2, 1e-2, 5e-3, 5e-3, 2.5e-3], rtol=1e-6)
    assert abs(m.current_learning_rate - 1e-2 * 0.5 ** 2) < 1e-12        # the logged rate follows the same staircase
    m.decay_learning_rate(0.1)
    m.single_step()
    assert abs(float(m.engine.sched[1]) - 2.5e-4) <= 1e-9
    # the decay factor survives a checkpoint round trip (it used to be lost: a resumed run got the undecayed rate)
    path = m.save_model_checkpoint(c.train_dir)
    m2 = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c)
    m2.load_model_from_checkpoint_file(path)
    assert abs(m2._lr_scale - 0.1) < 1e-15 and m2.current_step == 6 and int(m2.engine.step_dev[0]) == 6
    assert not [f for f in os.listdir(os.path.join(c.train_dir, "model")) if ".tmp" in f]


def test_bf16_fed_features_and_pinned_inputs_contract():
    """config.feature_dtype = 'bf16': the placeholder and the staging buffers are bf16; fp32 host batches are rounded
    on the host, bf16 CPU tensors are taken as they are; the maps equal those of a model fed the rounded fp32 features."""
    lib = EmuLib()
    c1, c2 = small_config(), small_config()
    c1.trainable = c2.trainable = False
    c2.feature_dtype = 'bf16'
    m1 = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c1)
    m2 = gaze_grcn.GazePredictionGRCN({'lib': lib}, data(), c2)
    assert m2.c3d_input.dtype == torch.bfloat16 and m1.c3d_input.dtype == torch.float32
    rng = np.random.RandomState(0)
    c3d = SY.make_c3d(rng, 4, 3, 64, 7)
    rounded = torch.from_numpy(c3d).to(torch.bfloat16)
    a = m1.predict(rounded.float().numpy())
    b = m2.predict(c3d)                      # fp32 batch, rounded by the model
    d = m2.predict(rounded)                  # bf16 CPU tensor
    assert np.array_equal(a, b) and np.array_equal(a, d)
    assert a.shape == (4, 3, 49, 49) and np.allclose(a.sum(axis=(2, 3)), 1.0, atol=1e-5)
