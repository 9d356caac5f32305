"""Host orchestration of the cascade (cascade_engine.py: 4-tap decomposition of the 11x11 / stride-7 transposed conv,
operand layout of the 5x5 top cell, tap tables, maxout-FC packing) over the NumPy emulation of the C-ABI against the
oracle's restatement of gaze_grcn_cascade.py:189-452 — CPU only."""
import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.cascade_engine import CascadeEngine
from emu_lib import EmuLib


@pytest.mark.parametrize("top_in", [64, 65])
def test_cascade_forward_matches_oracle(top_in):
    B, T, Ch, Cin, Cfeat = 2, 2, 16, 32, 64
    p = O.init_params_cascade(hidden=Ch, top_in=top_in, dim_proj=Cin, dim_feature=Cfeat, seed=3, regime="S",
                              fc_bias_scale=0.05)
    rng = np.random.RandomState(5)
    c3d = SY.make_c3d(rng, B, T, Cfeat, 7)
    gt, _ = SY.make_gazemaps(rng, B, T, 49, 49)
    sal = rng.uniform(size=(B, T, 49, 49)).astype(np.float32) if top_in == 65 else None
    eng = CascadeEngine(B=B, T=T, Ch=Ch, Cin=Cin, Cfeat=Cfeat, dtype=0, device="cpu", lib=EmuLib(), top_in=top_in,
                        keep_top_states=True)
    eng.load_tf_params(p)
    pred = eng.forward(c3d=torch.from_numpy(c3d), gt=torch.from_numpy(gt),
                       frm_sal=None if sal is None else torch.from_numpy(sal)).numpy()
    net = {}
    with torch.no_grad():
        ref = O.cascade_network(O.to_torch(p), torch.tensor(c3d, dtype=torch.float64),
                                frm_sal=None if sal is None else torch.tensor(sal, dtype=torch.float64), net=net)
        ref_loss = float(O.cascade_loss(ref, torch.tensor(gt, dtype=torch.float64)))
    for t in range(T):
        h_ref = net["rcn_outputs"][t].numpy()
        assert np.abs(eng.read_state(t + 1).numpy() - h_ref).max() <= 2e-5 * max(1e-6, np.abs(h_ref).max())
        u_ref = net["rcn_upsampled_outputs"][t].numpy()
        assert np.abs(eng.read_upsampled(t).numpy() - u_ref).max() <= 2e-5 * np.abs(u_ref).max(), t
        g_ref = net["top_outputs"][t].numpy()
        assert np.abs(eng.read_top_state(t).numpy() - g_ref).max() <= 2e-5 * np.abs(g_ref).max(), t
    ref = ref.numpy()
    assert ref.max() > 1e-3
    assert np.abs(pred - ref).max() <= 5e-5 * np.abs(ref).max()
    assert abs(float(eng.loss[0]) - ref_loss) <= 1e-5 * abs(ref_loss)
    # the zero borders of the top cell's operand layout are never written
    U = eng.U[:B * T * 2601].view(B * T, 51, 51, 64)
    assert not U[:, :2].any() and not U[:, :, :2].any()


def test_cascade_refuses_training_and_wrong_loss():
    with pytest.raises(NotImplementedError):
        CascadeEngine(B=1, T=1, Ch=16, Cin=32, Cfeat=64, dtype=0, device="cpu", lib=EmuLib(), train=True)
    eng = CascadeEngine(B=1, T=1, Ch=16, Cin=32, Cfeat=64, dtype=0, device="cpu", lib=EmuLib())
    with pytest.raises(NotImplementedError):
        eng.loss_fwd(torch.zeros(1, 1, 49, 49), "xentropy")


def test_cascade_param_roundtrip_and_model_class_contract(tmp_path):
    """load_tf_params / op_to_tf are inverses (checkpoints carry the reference's variable names and TF layouts); the
    model class mirrors gaze_grcn_cascade.py's builder signature and the GazePredictionGRU method contract."""
    from recurrent_gaze_prediction_b200.models import gaze_grcn_cascade as GC
    from recurrent_gaze_prediction_b200.models import gaze_rnn
    from recurrent_gaze_prediction_b200.models.train_gaze import get_model_class
    p = O.init_params_cascade(hidden=16, top_in=65, dim_proj=32, dim_feature=64, seed=1, regime="S", fc_bias_scale=0.1)
    eng = CascadeEngine(B=1, T=1, Ch=16, Cin=32, Cfeat=64, dtype=0, device="cpu", lib=EmuLib(), top_in=65)
    eng.load_tf_params(p)
    back = eng.op_to_tf(eng.params.numpy())
    assert sorted(back.keys()) == sorted(p.keys())
    for k, v in p.items():
        assert np.array_equal(back[k], np.asarray(v, dtype=np.float32)), k

    assert get_model_class("gaze_grcn_cascade")[0] is GC.GazePredictionGRCN
    c = gaze_rnn.GRUModelConfig()
    c.n_lstm_steps, c.batch_size, c.cascade_state_size = 2, 2, 16
    c.dim_feature, c.dim_cnn_proj, c.dtype, c.device = 64, 32, 'fp32', 'cpu'
    c.train_dir = str(tmp_path)
    ds = SY.read_synthetic_data_sets(6, 6, 6, T=2, dim_feature=64)
    with pytest.raises(NotImplementedError):
        GC.GazePredictionGRCN({'lib': EmuLib()}, ds, c)           # trainable defaults to True: refused
    c.trainable = False
    m = GC.GazePredictionGRCN({'lib': EmuLib()}, ds, c)
    assert m.train_op is None and m.config.loss_type == 'l2'
    assert tuple(m.predicted_gazemaps.shape) == (2, 2, 49, 49)
    assert m.single_step(train_mode=False) == 0 and np.isfinite(m.last_loss)
    ret = m.generate(ds.valid, max_instances=2)
    assert ret['pred_gazemap_list'].shape == (4, 49, 49)
    # checkpoint round trip through the reference's variable names
    m.save_model_checkpoint(str(tmp_path))
    m2 = GC.GazePredictionGRCN({'lib': EmuLib()}, ds, c)
    m2.engine.tparams.zero_()
    assert m2.reload_checkpoint()
    assert torch.equal(m2.engine.params, m.engine.params)
    # builder signature of the reference: (frame_images, c3d_input, gt_gazemap, dropout_keep_prob, net) -> (maps, loss, summaries)
    rng = np.random.RandomState(0)
    c3d = torch.from_numpy(SY.make_c3d(rng, 1, 2, 64, 7))
    gt = torch.from_numpy(SY.make_gazemaps(rng, 1, 2, 49, 49)[0])
    pp = O.init_params_cascade(hidden=16, top_in=64, dim_proj=32, dim_feature=64, seed=1, regime="S")
    net = {}
    pred, loss, summ = GC.GazePredictionGRCN.create_gazeprediction_network(None, c3d, gt, None, net, params=pp, lib=EmuLib())
    with torch.no_grad():
        ref = O.cascade_network(O.to_torch(pp), c3d.double())
        ref_loss = float(O.cascade_loss(ref, gt.double()))
    assert summ is None and np.abs(pred.numpy() - ref.numpy()).max() <= 5e-5 * float(ref.abs().max())
    assert abs(float(loss[0]) - ref_loss) <= 1e-5 * ref_loss and len(net['rcn_upsampled_outputs']) == 2
