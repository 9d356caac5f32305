"""Host orchestration (engine.py) over the NumPy emulation of the C-ABI vs the oracle — CPU only.
Checks the row layouts, tap tables, weight repacking, composed head and (below) the BPTT wiring."""
import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.engine import GazeEngine
from emu_lib import EmuLib


def _setup(model, B=2, T=3, Ch=16, G=7, seed=0, regime="S", entry="c3d", Cin=32, Cfeat=64, train=False, seq=None):
    p = O.init_params(model, hidden=Ch, T=T, grid=G, dim_proj=Cin, dim_feature=Cfeat, seed=seed, regime=regime)
    rng = np.random.RandomState(seed + 100)
    GH = 7 if model == "gaze_grcn77" else 49
    if model == "gaze_grcn77":
        GH = G
    c3d = SY.make_c3d(rng, B, T, Cfeat, G)
    emb = SY.make_embedded(rng, B, T, G, Cin)
    gt, _ = SY.make_gazemaps(rng, B, T, GH, GH, normalize=True)
    eng = GazeEngine(model, B=B, T=T, G=G, Ch=Ch, Cin=Cin, Cfeat=Cfeat, dtype=0, device="cpu", lib=EmuLib(),
                     train=train, seq=seq)
    eng.load_tf_params(p)
    return p, c3d, emb, gt, eng


@pytest.mark.parametrize("model", ["gaze_grcn", "gaze_grcn77", "gaze_lstm"])
@pytest.mark.parametrize("entry", ["c3d", "emb"])
def test_forward_matches_oracle(model, entry):
    p, c3d, emb, gt, eng = _setup(model)
    kw = dict(c3d=torch.from_numpy(c3d)) if entry == "c3d" else dict(emb=torch.from_numpy(emb))
    okw = dict(c3d=c3d.astype(np.float64)) if entry == "c3d" else dict(emb=emb.astype(np.float64))
    for lt in ("xentropy", "l2", "KLD"):
        logits = eng.forward(gt=torch.from_numpy(gt), loss_type=lt, **kw).numpy()
        ref_logits, ref_probs, ref_loss, _ = O.forward_loss_grads(model, p, gt.astype(np.float64), loss_type=lt, **okw)
        scale = np.abs(ref_logits).max()
        assert np.abs(logits - ref_logits).max() <= 2e-5 * scale
        if lt != "l2":
            assert np.allclose(eng.probs.numpy(), ref_probs, rtol=1e-4, atol=1e-9)
        assert abs(float(eng.loss[0]) - ref_loss) <= 1e-5 * abs(ref_loss) + 1e-7


def test_grid14_embedded_entry():
    p, c3d, emb, gt, eng = _setup("gaze_grcn77", B=1, T=2, Ch=8, G=14, entry="emb")
    logits = eng.forward(emb=torch.from_numpy(emb)).numpy()
    ref, _, _, _ = O.forward_loss_grads("gaze_grcn77", p, gt.astype(np.float64), emb=emb.astype(np.float64))
    assert np.abs(logits - ref).max() <= 2e-5 * np.abs(ref).max()


def test_param_roundtrip():
    p, *_, eng = _setup("gaze_lstm")
    back = eng.op_to_tf(eng.params.numpy())
    for k, v in p.items():
        assert np.allclose(back[k].reshape(np.asarray(v).shape), np.asarray(v, dtype=np.float32)), k


def _grad_check(model, entry, loss_type="xentropy", B=2, T=3, Ch=16, seq=None):
    p, c3d, emb, gt, eng = _setup(model, B=B, T=T, Ch=Ch, train=True, seq=seq)
    kw = dict(c3d=torch.from_numpy(c3d)) if entry == "c3d" else dict(emb=torch.from_numpy(emb))
    okw = dict(c3d=c3d.astype(np.float64)) if entry == "c3d" else dict(emb=emb.astype(np.float64))
    eng.forward(gt=torch.from_numpy(gt), loss_type=loss_type, **kw)
    eng.backward()
    got = eng.op_to_tf(eng.grads.numpy())
    _, _, _, ref = O.forward_loss_grads(model, p, gt.astype(np.float64), loss_type=loss_type, **okw)
    for k, r in ref.items():
        a = got[k].reshape(r.shape).astype(np.float64)
        scale = np.abs(r).max()
        # out_b's gradient is sum(softmax - gt) = 0 up to fp32 cancellation -> absolute floor
        assert np.abs(a - r).max() <= 2e-4 * scale + (1e-6 if k == "out_b" else 1e-12), (k, np.abs(a - r).max(), scale)
    return eng, p


@pytest.mark.parametrize("model", ["gaze_grcn", "gaze_grcn77", "gaze_lstm"])
@pytest.mark.parametrize("entry", ["c3d", "emb"])
def test_bptt_matches_oracle_autograd(model, entry):
    _grad_check(model, entry)


@pytest.mark.parametrize("model", ["gaze_grcn", "gaze_grcn77"])
def test_persistent_recurrence_wiring(model):
    """engine -> rgp_convgru_seq_fwd (one call for all T steps): same buffers as the per-step path, forward and BPTT."""
    p, c3d, emb, gt, eng = _setup(model, seq=True)
    assert eng.use_seq
    logits = eng.forward(c3d=torch.from_numpy(c3d), gt=torch.from_numpy(gt)).numpy()
    ref_logits, _, ref_loss, _ = O.forward_loss_grads(model, p, gt.astype(np.float64), c3d=c3d.astype(np.float64))
    assert np.abs(logits - ref_logits).max() <= 2e-5 * np.abs(ref_logits).max()
    assert not any(c[0] == "tapgemm" and c[1] in (2, 3) for c in eng.lib.calls[:0])
    _grad_check(model, "c3d", seq=True)


def test_bptt_l2_loss():
    _grad_check("gaze_grcn", "c3d", loss_type="l2")


def test_clip_adam_steps_match_oracle():
    """Two optimiser steps (global-norm clip + TF-Adam + staircase LR) against the oracle restatement."""
    model = "gaze_grcn"
    p, c3d, emb, gt, eng = _setup(model, train=True)
    names = list(p.keys())
    params = [np.asarray(p[k], dtype=np.float64).copy() for k in names]
    m = [np.zeros_like(x) for x in params]
    v = [np.zeros_like(x) for x in params]
    for step in range(2):
        eng.forward(c3d=torch.from_numpy(c3d), gt=torch.from_numpy(gt))
        eng.backward()
        eng.optimizer_step(lr0=1e-2, clip=1e-3)      # clip small enough to bite
        pd = dict(zip(names, params))
        _, _, _, grads = O.forward_loss_grads(model, pd, gt.astype(np.float64), c3d=c3d.astype(np.float64))
        gl = [grads[k] for k in names]
        gl, norm = O.clip_by_global_norm(gl, 1e-3)
        assert norm > 1e-3
        params, m, v = O.adam_step(params, gl, m, v, step + 1, O.learning_rate(step, 1e-2))
        got = eng.op_to_tf(eng.params.numpy())
        for k, r in zip(names, params):
            assert np.allclose(got[k].reshape(r.shape), r, rtol=2e-4, atol=1e-7), (step, k)


@pytest.mark.parametrize("G,B,head", [(7, 3, None), (14, 2, "fc")])
def test_persistent_convlstm_orchestration_matches_per_step(G, B, head):
    """engine wiring of rgp_convlstm_seq_fwd (buffers, counters zeroed per launch, training saves) == the per-step path"""
    outs = []
    for seq in (True, False):
        p = O.init_params("gaze_lstm", hidden=16, T=3, grid=G, dim_proj=32, dim_feature=64, seed=0, regime="S", head=head)
        rng = np.random.RandomState(5)
        emb = SY.make_embedded(rng, B, 3, G, 32)
        eng = GazeEngine("gaze_lstm", B=B, T=3, G=G, Ch=16, Cin=32, Cfeat=64, dtype=0, device="cpu", lib=EmuLib(),
                         train=True, seq=seq, head=head)
        assert eng.use_seq_lstm == seq
        eng.load_tf_params(p)
        for _ in range(2):       # twice: the counters are re-zeroed, the cell restarts from zero
            logits = eng.forward(emb=torch.from_numpy(emb)).numpy().copy()
        outs.append((logits, eng.Lsave.numpy().copy(), eng.read_state()[0].numpy().copy()))
    for x, y in zip(*outs):
        assert np.allclose(x, y, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("G,B,model", [(7, 3, "gaze_grcn"), (14, 2, "gaze_grcn77")])
def test_small_batch_persistent_convgru_orchestration_matches_per_step(G, B, model):
    """engine wiring of rgp_convgru_seq_small_fwd (counters zeroed per launch, BN offsets, training saves) == per-step path"""
    outs = []
    for seq in ("small", False):
        p = O.init_params(model, hidden=16, T=3, grid=G, dim_proj=32, dim_feature=64, seed=0, regime="S")
        emb = SY.make_embedded(np.random.RandomState(5), B, 3, G, 32)
        eng = GazeEngine(model, B=B, T=3, G=G, Ch=16, Cin=32, Cfeat=64, dtype=0, device="cpu", lib=EmuLib(), train=True, seq=seq)
        assert eng.use_seq_small == (seq == "small") and not eng.use_seq
        eng.load_tf_params(p)
        for _ in range(2):
            logits = eng.forward(emb=torch.from_numpy(emb)).numpy().copy()
        outs.append((logits, eng.Z.numpy().copy(), eng.Rs.numpy().copy(), eng.Cs.numpy().copy(), eng.read_state().numpy().copy()))
    for x, y in zip(*outs):
        assert np.allclose(x, y, rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("model", ["gaze_grcn", "gaze_grcn77"])
def test_small_batch_persistent_bptt_wiring(model):
    """engine -> rgp_convgru_seq_small_fwd + rgp_convgru_seq_small_bwd (one launch each for all T steps): gradients equal
    torch autograd on the oracle."""
    eng, _ = _grad_check(model, "c3d", seq="small", T=4)
    assert eng.use_seq_small


def test_small_batch_persistent_lstm_bptt_wiring():
    """engine -> rgp_convlstm_seq_fwd + rgp_convlstm_seq_small_bwd (one launch each): gradients equal torch autograd on the
    oracle, peephole gradients included."""
    eng, _ = _grad_check("gaze_lstm", "c3d", seq=True, T=4)
    assert eng.use_seq_lstm
    assert any(getattr(c, "__len__", None) and c[0] == "tapgemm" for c in eng.lib.calls)
