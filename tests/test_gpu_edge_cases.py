"""Edge shapes of the round-2 kernels on the GPU: single clip / single step / ragged tiles, against the float64 oracle."""
import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.cascade_engine import CascadeEngine
from recurrent_gaze_prediction_b200.dense_engine import DenseGruEngine
from recurrent_gaze_prediction_b200.engine import GazeEngine

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("G,B,T", [(7, 1, 1), (14, 1, 2), (7, 3, 1), (3, 5, 4)])
def test_persistent_convlstm_smallest_shapes(G, B, T):
    """one clip / one step: a single, mostly out-of-range tile; G = 3: several frames per tile"""
    p = O.init_params("gaze_lstm", hidden=128, T=T, grid=G, seed=1, regime="S", head="fc")
    emb = SY.make_embedded(np.random.RandomState(2), B, T, G, 512)
    eng = GazeEngine("gaze_lstm", B=B, T=T, G=G, Ch=128, dtype=1, device="cuda", head="fc")
    assert eng.use_seq_lstm
    eng.load_tf_params(p)
    eng.forward(emb=torch.from_numpy(emb).cuda())
    torch.cuda.synchronize()
    assert not eng.seq_timed_out()
    net = {}
    with torch.no_grad():
        O.recurrent_network("gaze_lstm", O.to_torch(p), emb=torch.tensor(emb, dtype=torch.float64), net=net)
    for t in range(T):
        h_ref = net["rcn_outputs"][t].numpy()
        got = eng.read_state(t + 1)[1].cpu().numpy()
        assert np.abs(got - h_ref).max() <= 4e-2 * max(1.0, np.abs(h_ref).max()), t


def test_i2h_dense_mode_single_frame_and_ragged_tail():
    """dense-M i2h (TMA im2col): 1 frame (49 rows: one partial tile) and 131 frames (6419 rows: last tile 19 rows)"""
    for B, T in ((1, 1), (131, 1)):
        p = O.init_params("gaze_grcn77", hidden=128, T=T, seed=4, regime="S")
        emb = SY.make_embedded(np.random.RandomState(3), B, T, 7, 512)
        eng = GazeEngine("gaze_grcn77", B=B, T=T, Ch=128, dtype=1, device="cuda")
        eng.load_tf_params(p)
        eng.stage_embed(emb=torch.from_numpy(emb).cuda())
        eng.stage_i2h()
        torch.cuda.synchronize()
        xg = eng.Xg.view(T * B, 8, 8, -1)[:, :7, :7].cpu().numpy()
        pt = O.to_torch(p)
        x = torch.tensor(emb[:, 0], dtype=torch.float64)
        ref = torch.cat([O.conv2d_same(x, pt["GRU_Conv_" + k]) for k in ("Wz", "Wr", "W")], dim=3).numpy()
        assert np.abs(xg - ref).max() <= 2e-2 * np.abs(ref).max()


def test_cascade_single_clip_single_step():
    p = O.init_params_cascade(seed=2, regime="S", fc_bias_scale=0.05)
    c3d = SY.make_c3d(np.random.RandomState(1), 1, 1)
    for dtype, tol in ((0, 1e-4), (1, 2e-2)):
        eng = CascadeEngine(B=1, T=1, dtype=dtype, device="cuda")
        eng.load_tf_params(p)
        pred = eng.forward(c3d=torch.from_numpy(c3d).cuda()).cpu().numpy()
        with torch.no_grad():
            ref = O.cascade_network(O.to_torch(p), torch.tensor(c3d, dtype=torch.float64)).numpy()
        assert np.abs(pred - ref).max() <= tol * ref.max()


def test_dense_gru_training_single_clip():
    p = O.init_params_dense(seed=5)
    rng = np.random.RandomState(2)
    c3d = SY.make_c3d(rng, 1, 2)
    gt = O.normalize_probability_map(SY.make_gazemaps(rng, 1, 2, 49, 49)[0]).astype(np.float32)
    e = DenseGruEngine(B=1, T=2, dtype=0, train=True)
    e.load_tf_params(p)
    e.forward(c3d=torch.from_numpy(c3d).cuda(), gt=torch.from_numpy(gt).cuda())
    e.backward()
    got = e.op_to_tf(e.grads.cpu().numpy())
    _, _, ref_loss, ref = O.forward_loss_grads("gaze_rnn", p, gt.astype(np.float64), c3d=c3d.astype(np.float64))
    assert abs(float(e.loss.cpu()[0]) - ref_loss) <= 1e-5 * abs(ref_loss)
    for k, r in ref.items():
        assert np.abs(got[k].reshape(r.shape) - r).max() <= 1e-3 * max(np.abs(r).max(), 1e-30), k
