"""Reference-facing model classes and entry points on the real library (B200)."""
import os

import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.models import evaluate_gaze, gaze_grcn, gaze_lstm, gaze_rnn, train_gaze

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype,tol", [("fp32", 1e-4), ("bf16", 2e-2)])
def test_network_builder_matches_oracle(dtype, tol):
    rng = np.random.RandomState(0)
    B, T = 2, 4
    c3d = SY.make_c3d(rng, B, T)
    p = O.init_params("gaze_grcn", hidden=128, T=T, seed=1, regime="S")
    net = {}
    logits = gaze_grcn.GazePredictionGRCN.create_gazeprediction_network(None, torch.from_numpy(c3d).cuda(), net=net,
                                                                        params=p, dtype=dtype)
    gt = np.full((B, T, 49, 49), 1.0 / 2401)
    ref, ref_probs, _, _ = O.forward_loss_grads("gaze_grcn", p, gt, c3d=c3d.astype(np.float64))
    probs = torch.softmax(logits.reshape(B, T, -1).double(), -1).reshape(B, T, 49, 49).cpu().numpy()
    rel = np.abs(probs - ref_probs).max(axis=(2, 3)) / ref_probs.max(axis=(2, 3))
    assert rel.max() <= tol
    assert tuple(net['c3d_embedded'].shape) == (B, T, 7, 7, 512)


@pytest.mark.parametrize("dtype,tol", [("fp32", 2e-5), ("bf16", 2e-2)])
def test_cells_on_gpu(dtype, tol):
    rng = np.random.RandomState(3)
    B, G, Cin, Ch = 3, 7, 512, 128
    x = torch.from_numpy(rng.standard_normal((B, G, G, Cin)).astype(np.float32)).cuda()
    cell = gaze_grcn.GRU_RCN_Cell(Ch, Cin, dtype=dtype)
    big = {k: v * 300 for k, v in cell.variables.items()}
    cell.load_variables(big)
    h0 = torch.from_numpy(rng.standard_normal((B, G, G, Ch)).astype(np.float32)).cuda()
    _, h1 = cell(x, h0)
    _, ref = O.gru_rcn_cell(x.cpu().double(), h0.cpu().double(), {k: torch.tensor(v) for k, v in big.items()})
    assert (h1.cpu().double() - ref).abs().max() <= tol * max(1.0, ref.abs().max())
    lc = gaze_lstm.LSTM_RCN_Cell(Ch, Cin, dtype=dtype)
    bigl = {k: v * 300 for k, v in lc.variables.items()}
    lc.load_variables(bigl)
    s0 = torch.from_numpy(rng.standard_normal((B, G, G, 2 * Ch)).astype(np.float32)).cuda()
    h, s1 = lc(x, s0)
    refh, refs = O.lstm_rcn_cell(x.cpu().double(), s0.cpu().double(), {k: torch.tensor(v) for k, v in bigl.items()})
    assert (s1.cpu().double() - refs).abs().max() <= tol * max(1.0, refs.abs().max())


def test_train_and_evaluate_entry_points(tmp_path):
    args = train_gaze.make_parser().parse_args(
        ["--model", "gaze_grcn", "--dataset", "synthetic", "--max_steps", "6", "--batch_size", "4", "--n_lstm_steps", "5",
         "--train_dir", str(tmp_path), "--learning_rate", "0.001", "--dtype", "bf16"])
    np.random.seed(0)
    model, scores = train_gaze.train(args)
    assert model.current_step == 7 and np.isfinite(model.last_loss)
    assert sorted(scores) == ['AUC_Borji', 'AUC_shuffled', 'cc', 'sim']
    path = model.save_model_checkpoint(str(tmp_path))
    out = evaluate_gaze.main(["--model", "gaze_grcn", "--dataset", "synthetic", "--checkpoint_path", path,
                              "--num_frames", "30", "--batch_size", "4", "--n_lstm_steps", "5",
                              "--fixation_original_scale"])
    assert len(out['AUC_Judd']) == 40 and np.isfinite(np.nanmean(out['AUC_Judd']))
    gen = os.path.join(str(tmp_path), "generated-synthetic-7", "overall.txt")
    assert os.path.isfile(gen)


@pytest.mark.parametrize("mod,cls", [(gaze_grcn, "GazePredictionGRCN"), (gaze_lstm, "GazePredictionLSTM")])
def test_training_through_the_model_class_learns(mod, cls):
    """30 Adam steps on one repeated synthetic batch: the loss must fall (end-to-end sanity of fwd+BPTT+Adam)."""
    c = gaze_rnn.GRUModelConfig()
    c.n_lstm_steps, c.batch_size, c.initial_learning_rate, c.use_flip_batch = 6, 4, 1e-3, False
    ds = SY.read_synthetic_data_sets(4, 4, 4, T=6)
    m = getattr(mod, cls)(None, ds, c)
    losses = []
    for _ in range(30):
        m.single_step()
        losses.append(m.last_loss)
    assert losses[-1] < losses[0] - 0.05, (losses[0], losses[-1])


@pytest.mark.parametrize("B,hidden", [(2, 128), (256, 256)])
def test_long_clip_extraction_chunked_equals_one_pass(tmp_path, B, hidden):
    """extract_map.py (n_lstm_steps = 105): three 35-frame chunks with the recurrent state carried between them against
    one 105-frame pass.  Same kernels on the same numbers -> identical bits.  B = 256 takes the persistent recurrence
    kernel (its stateful chunking), B = 2 the per-step GEMMs."""
    from recurrent_gaze_prediction_b200.models import extract_map as XM
    T = 105 if B == 2 else 15
    Tc = 35 if B == 2 else 5
    argv = ["--model", "gaze_grcn", "--dataset", "synthetic", "--batch_size", str(B), "--n_lstm_steps", str(T),
            "--hidden", str(hidden), "--num_clips", str(B), "--out_dir", str(tmp_path)]
    args1 = XM.make_parser().parse_args(argv + ["--chunk_steps", "0"])
    args2 = XM.make_parser().parse_args(argv + ["--chunk_steps", str(Tc)])
    one = XM.load_model(args1)
    chk = XM.load_model(args2)
    assert one.chunk_steps is None and chk.chunk_steps == Tc and chk.engine.T == Tc and chk.engine.Tp == T
    assert chk.engine.use_seq == (B == 256)
    g = one.engine.view("bn_gamma")
    g.copy_(torch.linspace(0.5, 1.5, g.numel(), device=g.device).view_as(g))
    chk.engine.params.copy_(one.engine.params)
    one.engine.refresh_operands()
    chk.engine.refresh_operands()
    clips = list(XM.iter_clips(args1))
    r1 = XM.extract(one, clips)
    r2 = XM.extract(chk, clips, out_dir=str(tmp_path))
    assert len(r1) == B
    for name, c3d in clips[:4] + clips[-2:]:
        n = min(T, len(c3d))
        assert r1[name][0].shape == (n, 49, 49) and r1[name][1].shape == (n, 7, 7)
        assert np.array_equal(r1[name][0], r2[name][0]), name
        assert np.allclose(r1[name][0].sum(axis=(1, 2)), 1.0, atol=1e-4)
    assert os.path.isfile(os.path.join(str(tmp_path), clips[0][0], clips[0][0] + ".gazemap.npy"))


def test_device_flip_matches_host_flip():
    """gaze_rnn.py:504-510: the W-axis flip of the chosen clips, done on the fed device buffers (rgp_flip_w)."""
    from recurrent_gaze_prediction_b200 import lib as L
    lib = L.get_lib()
    rng = np.random.RandomState(0)
    for shape in ((6, 5, 1024, 7, 7), (6, 5, 49, 49), (3, 2, 4, 1), (5, 3, 8)):
        x = rng.standard_normal(shape).astype(np.float32)
        sel = np.array([i for i in (4, 1, 2) if i < shape[0]], dtype=np.int32)
        want = x.copy()
        want[sel] = want[sel][..., ::-1]
        d = torch.from_numpy(x).cuda()
        idx = torch.from_numpy(sel).cuda()
        lib.rgp_flip_w(d.data_ptr(), idx.data_ptr(), len(sel), int(np.prod(shape[1:-1])), shape[-1],
                       torch.cuda.current_stream().cuda_stream)
        assert np.array_equal(d.cpu().numpy(), want), shape
    lib.rgp_flip_w(d.data_ptr(), idx.data_ptr(), 0, 1, 1, 0)          # nothing selected: no-op


@pytest.mark.parametrize("dtype,tol,B", [("fp32", 1e-4, 3), ("bf16", 2e-2, 3), ("bf16", 2e-2, 130)])
def test_dense_gru_variant_on_gpu(dtype, tol, B):
    """SURVEY row a24 at the reference's sizes (1617 units, proj_out to 49x49): gaze maps against the oracle's
    restatement of the tf GRUCell (fp32 1e-4 / bf16 2e-2 relative to the map's peak, as for the GRCN path)."""
    from recurrent_gaze_prediction_b200.dense_engine import DenseGruEngine
    rng = np.random.RandomState(1)
    T = 4
    p = O.init_params_dense(seed=3)
    e = DenseGruEngine(B=B, T=T, dtype=0 if dtype == "fp32" else 1)
    assert (e.H, e.Ip, e.Hp, e.Np) == (1617, 1600, 1664, 2432)
    e.load_tf_params(p)
    c3d = SY.make_c3d(rng, B, T)
    gt = O.normalize_probability_map(SY.make_gazemaps(rng, B, T, 49, 49)[0]).astype(np.float32)
    e.forward(c3d=torch.from_numpy(c3d).cuda(), gt=torch.from_numpy(gt).cuda())
    nb = min(B, 4)                                             # the oracle on the first clips (clips are independent)
    _, ref_probs, _, _ = O.forward_loss_grads("gaze_rnn", p, gt[:nb].astype(np.float64), c3d=c3d[:nb].astype(np.float64))
    probs = e.probs.cpu().numpy()[:nb]
    rel = np.abs(probs - ref_probs).max(axis=(2, 3)) / ref_probs.max(axis=(2, 3))
    assert rel.max() <= tol, rel.max()
    assert np.isfinite(float(e.loss.cpu()[0]))
    assert float(e.Hf[:, e.H:].abs().max()) == 0.0
    if B == 130:                                               # last clips too (second M tile of the per-step GEMMs)
        _, ref2, _, _ = O.forward_loss_grads("gaze_rnn", p, gt[-2:].astype(np.float64), c3d=c3d[-2:].astype(np.float64))
        rel = np.abs(e.probs.cpu().numpy()[-2:] - ref2).max(axis=(2, 3)) / ref2.max(axis=(2, 3))
        assert rel.max() <= tol, rel.max()


@pytest.mark.parametrize("dtype,tol", [("fp32", 1e-3), ("bf16", 5e-2)])
def test_dense_gru_training_on_gpu(dtype, tol):
    """SURVEY row a24, training half (gaze_rnn.py:447-478): hand-written BPTT of the dense variant at the reference's
    sizes against torch autograd on the float64 oracle (per tensor, relative to its largest gradient: fp32 1e-3, bf16
    5e-2 - the bars of the GRCN BPTT tests), then clip + Adam steps through the model class with live dropout."""
    from recurrent_gaze_prediction_b200.dense_engine import DenseGruEngine
    from recurrent_gaze_prediction_b200.models import gaze_rnn
    rng = np.random.RandomState(2)
    B, T = 3, 4
    p = O.init_params_dense(seed=3)
    e = DenseGruEngine(B=B, T=T, dtype=0 if dtype == "fp32" else 1, train=True)
    e.load_tf_params(p)
    c3d = SY.make_c3d(rng, B, T)
    gt = O.normalize_probability_map(SY.make_gazemaps(rng, B, T, 49, 49)[0]).astype(np.float32)
    e.forward(c3d=torch.from_numpy(c3d).cuda(), gt=torch.from_numpy(gt).cuda())
    e.backward()
    torch.cuda.synchronize()
    got = e.op_to_tf(e.grads.cpu().numpy())
    _, _, ref_loss, ref = O.forward_loss_grads("gaze_rnn", p, gt.astype(np.float64), c3d=c3d.astype(np.float64))
    assert abs(float(e.loss.cpu()[0]) - ref_loss) <= (1e-5 if dtype == "fp32" else 2e-2) * abs(ref_loss)
    for k, r in ref.items():
        a = got[k].reshape(r.shape).astype(np.float64)
        assert np.abs(a - r).max() <= tol * max(np.abs(r).max(), 1e-30), (k, np.abs(a - r).max(), np.abs(r).max())
    assert float(e.grads[torch.from_numpy(e.pad_mask() == 0).cuda()].abs().max()) == 0.0
    # dropout: forward mask == backward mask (the gradient of the dropped inputs' projection rows is exactly zero)
    e.keep_prob, e.dropout_seed = 0.5, 77
    e.forward(c3d=torch.from_numpy(c3d).cuda(), gt=torch.from_numpy(gt).cuda())
    e.backward()
    X = e.X[:T * B, :e.I].float().cpu().numpy()
    dP = e.dPacc[:T * B * e.P].float().cpu().numpy().reshape(T * B, e.I)
    frac = (X != 0).mean()
    assert 0.45 < frac < 0.55 and not dP[X == 0].any() and np.abs(dP[X != 0]).max() > 0
    # the model class trains it
    c = gaze_rnn.GRUModelConfig()
    c.n_lstm_steps, c.batch_size, c.dtype, c.initial_learning_rate, c.steps_per_logprint = T, 2, dtype, 1e-3, 1000
    ds = SY.read_synthetic_data_sets(6, 6, 6, T=T)
    m = gaze_rnn.GazePredictionGRU(None, ds, c)
    p0 = m.engine.params.clone()
    losses = []
    for k in range(3):
        assert m.single_step(train_mode=True) == k + 1
        losses.append(m.last_loss)
    assert all(np.isfinite(losses)) and float((m.engine.params - p0).abs().max()) > 1e-4
