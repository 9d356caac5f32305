"""GPU parity tests proper: the CUDA path (through the C-ABI) against the oracle on seeded inputs.
Tolerances are the north star's: fp32 gaze maps <= 1e-4 relative, bf16 tensor-core path <= 2e-2."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import lib as L
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.engine import GazeEngine
from emu_lib import EmuLib

pytestmark = pytest.mark.gpu
TOL = {0: 1e-4, 1: 2e-2}


def _dev_ptr(t):
    return t.data_ptr()


def _run_tapgemm_case(dtype, engine, M, N, K, taps, seed, epi=L.EPI_STORE_F32):
    """Generic tap GEMM (STORE_F32) on the GPU vs the NumPy emulation on the same (rounded) operands."""
    rng = np.random.RandomState(seed)
    rowsA = M + max(taps) + 5
    tdt = torch.float32 if dtype == 0 else torch.bfloat16
    A = torch.from_numpy(rng.standard_normal((rowsA, K)).astype(np.float32)).to(tdt)
    Bw = torch.from_numpy((rng.standard_normal((len(taps) * N, K)) / np.sqrt(K * len(taps))).astype(np.float32)).to(tdt)
    Af, Bf = A.float().contiguous(), Bw.float().contiguous()
    out_ref = torch.zeros(M, N)

    def args(Ap, Bp, outp, dt):
        a = L.TapGemmArgs()
        a.A, a.rowsA, a.lda, a.Bw = Ap, rowsA, K, Bp
        a.M, a.N, a.K, a.ntaps = M, N, K, len(taps)
        for i, o in enumerate(taps):
            a.tap_off[i] = o
        a.dtype, a.engine, a.epi = dt, engine, epi
        a.out_f32, a.ldo = outp, N
        return a

    EmuLib().rgp_tapgemm(C.byref(args(Af.data_ptr(), Bf.data_ptr(), out_ref.data_ptr(), 0)), 0)
    Ad, Bd = A.cuda(), Bw.cuda()
    out = torch.zeros(M, N, device="cuda")
    L.get_lib().rgp_tapgemm(C.byref(args(Ad.data_ptr(), Bd.data_ptr(), out.data_ptr(), dtype)),
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    err = (out.cpu() - out_ref).abs().max().item()
    return err, out_ref.abs().max().item()


@pytest.mark.parametrize("dtype,engine", [(0, 0), (1, 1), (1, 0)])
@pytest.mark.parametrize("M,N,K,taps", [
    (300, 96, 64, [0]),
    (1000, 768, 512, [0, 1, 2, 8, 9, 10, 16, 17, 18]),
    (515, 48, 128, [33 - 10 * dy - dx for dy in range(4) for dx in range(4)]),
    (128, 256, 1024, [0]),
    (4096, 512, 256, [0, 1, 2, 8, 9, 10, 16, 17, 18]),
])
def test_tapgemm_matches_emulation(dtype, engine, M, N, K, taps):
    err, scale = _run_tapgemm_case(dtype, engine, M, N, K, taps, seed=M + N)
    assert err <= 2e-5 * max(scale, 1.0) * (1 if dtype == 0 else 4), (err, scale)


def _case(model, B, T, Ch, G=7, seed=0, regime="S", entry="c3d", Cin=512, Cfeat=1024):
    p = O.init_params(model, hidden=Ch, T=T, grid=G, dim_proj=Cin, dim_feature=Cfeat, seed=seed, regime=regime)
    rng = np.random.RandomState(seed + 100)
    GH = G if model == "gaze_grcn77" else 49
    c3d = SY.make_c3d(rng, B, T, Cfeat, G) if entry == "c3d" else None
    emb = SY.make_embedded(rng, B, T, G, Cin) if entry == "emb" else None
    gt, _ = SY.make_gazemaps(rng, B, T, GH, GH, normalize=True)
    return p, c3d, emb, gt


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("model,Ch,regime,entry", [
    ("gaze_grcn", 128, "S", "c3d"),
    ("gaze_grcn", 256, "S", "c3d"),
    ("gaze_grcn", 128, "R", "c3d"),
    ("gaze_grcn77", 128, "S", "c3d"),
    ("gaze_lstm", 128, "S", "c3d"),
    ("gaze_grcn", 128, "S", "emb"),
])
def test_forward_parity(dtype, model, Ch, regime, entry):
    B, T = 3, 6
    p, c3d, emb, gt = _case(model, B, T, Ch, regime=regime, entry=entry)
    okw = dict(c3d=c3d.astype(np.float64)) if entry == "c3d" else dict(emb=emb.astype(np.float64))
    ref_logits, ref_probs, ref_loss, _ = O.forward_loss_grads(model, p, gt.astype(np.float64), **okw)
    eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=dtype, device="cuda")
    eng.load_tf_params(p)
    kw = dict(c3d=torch.from_numpy(c3d).cuda()) if entry == "c3d" else dict(emb=torch.from_numpy(emb).cuda())
    eng.forward(gt=torch.from_numpy(gt).cuda(), **kw)
    torch.cuda.synchronize()
    probs = eng.probs.cpu().numpy().astype(np.float64)
    logits = eng.logits.cpu().numpy().astype(np.float64)
    tol = TOL[dtype]
    # gaze maps: relative to the map's peak (the quantity the saliency metrics consume)
    rel = np.abs(probs - ref_probs).max(axis=(2, 3)) / ref_probs.max(axis=(2, 3))
    assert rel.max() <= tol, (rel.max(), np.abs(logits - ref_logits).max())
    assert abs(float(eng.loss.cpu()[0]) - ref_loss) <= tol * abs(ref_loss)


@pytest.mark.parametrize("dtype", [0, 1])
def test_grid14_long_sequence_lstm_and_gru(dtype):
    """BASELINE config 4 shape family: 14x14 grid, embedded-feature entry, longer T (kept small here)."""
    for model in ("gaze_grcn77",):
        B, T, Ch, G = 2, 12, 128, 14
        p, _, emb, gt = _case(model, B, T, Ch, G=G, entry="emb")
        ref_logits, ref_probs, ref_loss, _ = O.forward_loss_grads(model, p, gt.astype(np.float64),
                                                                  emb=emb.astype(np.float64))
        eng = GazeEngine(model, B=B, T=T, G=G, Ch=Ch, dtype=dtype, device="cuda")
        eng.load_tf_params(p)
        eng.forward(emb=torch.from_numpy(emb).cuda(), gt=torch.from_numpy(gt).cuda())
        torch.cuda.synchronize()
        probs = eng.probs.cpu().numpy().astype(np.float64)
        rel = np.abs(probs - ref_probs).max(axis=(2, 3)) / ref_probs.max(axis=(2, 3))
        assert rel.max() <= TOL[dtype], rel.max()


@pytest.mark.parametrize("loss_type", ["xentropy", "l2", "KLD"])
def test_softmax_loss_kernel(loss_type):
    rng = np.random.RandomState(3)
    n, hw = 37, 2401
    logits = (rng.standard_normal((n, hw)) * 3).astype(np.float32)
    gt = rng.uniform(size=(n, hw)).astype(np.float32)
    gt /= gt.sum(1, keepdims=True)
    emu = EmuLib()
    pr, dl, fl = np.zeros((n, hw), np.float32), np.zeros((n, hw), np.float32), np.zeros(n, np.float32)
    lt = {"xentropy": 0, "l2": 1, "KLD": 2}[loss_type]
    emu.rgp_softmax_xent_fwd_bwd(logits.ctypes.data, gt.ctypes.data, pr.ctypes.data, dl.ctypes.data, fl.ctypes.data,
                                 n, hw, 1.0 / n, lt, 0)
    lg, g = torch.from_numpy(logits).cuda(), torch.from_numpy(gt).cuda()
    P, D, F = torch.zeros_like(lg), torch.zeros_like(lg), torch.zeros(n, device="cuda")
    L.get_lib().rgp_softmax_xent_fwd_bwd(lg.data_ptr(), g.data_ptr(), P.data_ptr(), D.data_ptr(), F.data_ptr(),
                                         n, hw, 1.0 / n, lt, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.allclose(P.cpu().numpy(), pr, rtol=2e-5, atol=1e-9)
    assert np.allclose(D.cpu().numpy(), dl, rtol=2e-4, atol=1e-9)
    assert np.allclose(F.cpu().numpy(), fl, rtol=2e-5)


def test_library_is_the_cuda_one():
    lib = L.get_lib()
    assert lib.path.endswith("librgp_sm100.so") and lib.has_tcgen05()


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("model,Ch,entry", [("gaze_grcn", 128, "c3d"), ("gaze_grcn77", 128, "emb"),
                                            ("gaze_grcn", 256, "c3d"), ("gaze_lstm", 128, "c3d"),
                                            ("gaze_lstm", 256, "emb")])
def test_bptt_parity(dtype, model, Ch, entry):
    """Hand-written BPTT vs torch autograd on the oracle (float64).  fp32 path: 1e-3 of each tensor's max;
    bf16 path: 5e-2 (operands and saved activations are rounded to bf16)."""
    B, T = 3, 5
    p, c3d, emb, gt = _case(model, B, T, Ch, regime="S", entry=entry)
    okw = dict(c3d=c3d.astype(np.float64)) if entry == "c3d" else dict(emb=emb.astype(np.float64))
    _, _, ref_loss, ref = O.forward_loss_grads(model, p, gt.astype(np.float64), **okw)
    eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=dtype, device="cuda", train=True)
    eng.load_tf_params(p)
    kw = dict(c3d=torch.from_numpy(c3d).cuda()) if entry == "c3d" else dict(emb=torch.from_numpy(emb).cuda())
    eng.forward(gt=torch.from_numpy(gt).cuda(), **kw)
    eng.backward()
    torch.cuda.synchronize()
    got = eng.op_to_tf(eng.grads.cpu().numpy())
    tol = 1e-3 if dtype == 0 else 5e-2
    for k, r in ref.items():
        if entry == "emb" and k.startswith("proj_c3d"):
            continue
        a = got[k].reshape(r.shape).astype(np.float64)
        scale = np.abs(r).max()
        err = np.abs(a - r).max()
        assert err <= tol * scale + (1e-6 if k == "out_b" else 1e-12), (k, err, scale)


@pytest.mark.parametrize("dtype", [0, 1])
@pytest.mark.parametrize("model", ["gaze_grcn", "gaze_lstm"])
def test_training_steps_reduce_loss_and_track_oracle(dtype, model):
    """3 optimiser steps (clip + TF-Adam) on the GPU vs the oracle restatement (fp32: parameters agree)."""
    B, T, Ch = 2, 4, 128
    p, c3d, emb, gt = _case(model, B, T, Ch, regime="S")
    eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=dtype, device="cuda", train=True)
    eng.load_tf_params(p)
    names = list(p.keys())
    params = [np.asarray(p[k], dtype=np.float64).copy() for k in names]
    m = [np.zeros_like(x) for x in params]
    v = [np.zeros_like(x) for x in params]
    c3d_d, gt_d = torch.from_numpy(c3d).cuda(), torch.from_numpy(gt).cuda()
    losses = []
    for step in range(3):
        eng.forward(c3d=c3d_d, gt=gt_d)
        losses.append(float(eng.loss.cpu()[0]))
        eng.backward()
        eng.optimizer_step(lr0=1e-3, clip=10.0)
        _, _, ref_loss, grads = O.forward_loss_grads(model, dict(zip(names, params)), gt.astype(np.float64),
                                                     c3d=c3d.astype(np.float64))
        assert abs(losses[-1] - ref_loss) <= (1e-4 if dtype == 0 else 2e-2) * abs(ref_loss)
        gl, _ = O.clip_by_global_norm([grads[k] for k in names], 10.0)
        params, m, v = O.adam_step(params, gl, m, v, step + 1, O.learning_rate(step, 1e-3))
    assert losses[-1] < losses[0]
    if dtype == 0:
        got = eng.op_to_tf(eng.params.cpu().numpy())
        for k, r in zip(names, params):
            if k in ("out_b", "ConvLSTM_Whc"):
                continue   # (ConvLSTM_Whc never gets a gradient.)  out_b: its true gradient is sum(softmax - gt) == 0; Adam normalises fp32 rounding noise to ~lr
            # Adam's first steps move every weight by ~lr regardless of gradient size: compare the displacement
            d_ref = r - np.asarray(p[k], dtype=np.float64)
            d_got = got[k].reshape(r.shape) - np.asarray(p[k], dtype=np.float32)
            assert np.abs(d_got - d_ref).max() <= 0.05 * np.abs(d_ref).max() + 1e-7, k


@pytest.mark.parametrize("engine", [0, 1])
@pytest.mark.parametrize("M,Ny,Nx,ldy,ycol0,taps", [
    (960, 256, 128, 384, 0, [0, 1, 2, 8, 9, 10, 16, 17, 18]),      # U_zr wgrad shape at Ch=128
    (1000, 128, 128, 384, 256, [0, 1, 2, 8, 9, 10, 16, 17, 18]),   # U_c: column window of dG, M % 64 != 0
    (20000, 768, 512, 768, 0, [0, 9, 18]),                          # W_x shape: several row slabs per tile (stream-K)
    (4100, 384, 256, 384, 0, [0]),                                  # composed-head wgrad
    (3000, 512, 1024, 512, 0, [0]),                                 # projection wgrad
    (700, 200, 64, 200, 0, [0, 3]),                                 # ragged Ny (not a multiple of 128), BN = 64
])
def test_wgrad_matches_numpy(engine, M, Ny, Nx, ldy, ycol0, taps):
    """Weight-gradient contraction over rows (tcgen05 MN-major engine, engine=0; CUDA-core engine, engine=1)."""
    rng = np.random.RandomState(M + Ny)
    rowsX = M + max(taps) + 7
    Y = torch.from_numpy(rng.standard_normal((M, ldy)).astype(np.float32)).to(torch.bfloat16)
    X = torch.from_numpy(rng.standard_normal((rowsX, Nx)).astype(np.float32)).to(torch.bfloat16)
    Yf, Xf = Y.float().numpy().astype(np.float64), X.float().numpy().astype(np.float64)
    ref = np.stack([Yf[:, ycol0:ycol0 + Ny].T @ Xf[o:o + M] for o in taps])
    Yd, Xd = Y.cuda(), X.cuda()
    D = torch.zeros(len(taps), Ny, Nx, device="cuda")
    a = L.WgradArgs()
    a.Y, a.ldy, a.ycol0, a.Ny = Yd.data_ptr(), ldy, ycol0, Ny
    a.X, a.rowsX, a.ldx, a.xcol0, a.Nx = Xd.data_ptr(), rowsX, Nx, 0, Nx
    a.M, a.ntaps = M, len(taps)
    for i, o in enumerate(taps):
        a.tap_off[i] = o
    a.D, a.dtype, a.engine = D.data_ptr(), 1, engine
    lib = L.get_lib()
    lib.rgp_wgrad(C.byref(a), torch.cuda.current_stream().cuda_stream)
    lib.rgp_wgrad(C.byref(a), torch.cuda.current_stream().cuda_stream)      # accumulates: D = 2 * contraction
    torch.cuda.synchronize()
    got = D.cpu().numpy().astype(np.float64) / 2
    err = np.abs(got - ref).max()
    assert err <= 2e-5 * np.sqrt(M) * 4, (err, np.abs(ref).max())


@pytest.mark.parametrize("model,Ch,B,nsplit,mcast,train,ntile", [
    ("gaze_grcn", 256, 3, 1, 1, False, 1),
    ("gaze_grcn", 256, 6, 2, 1, False, 1),
    ("gaze_grcn", 128, 5, 1, 2, True, 1),      # 3 tiles in clusters of 2: one ghost tile
    ("gaze_grcn", 256, 4, 2, 2, True, 1),
    ("gaze_grcn", 256, 14, 2, 4, True, 1),     # 8-CTA clusters: 4 tiles x 2 channel halves, one ghost tile
    ("gaze_grcn", 256, 18, 1, 8, False, 1),
    ("gaze_grcn77", 128, 2, 1, 4, False, 1),
    ("gaze_grcn", 256, 320, 1, 4, False, 1),   # more tiles than SMs: CTAs walk several tiles
    ("gaze_grcn", 256, 8, 2, 1, False, 2),     # two tiles per CTA (every weight box feeds both), channel split 2
    ("gaze_grcn", 256, 5, 2, 1, True, 2),      # odd number of tiles: the second tile of the last pair is half / out of range
    ("gaze_grcn", 256, 2, 2, 1, True, 2),      # a single tile: the pair's second tile is a ghost
    ("gaze_grcn", 256, 600, 2, 1, False, 2),   # 150 pairs on 74 clusters: CTAs walk several pairs
    ("gaze_grcn", 256, 6, 4, 1, False, 1),     # four CTAs per tile (64 channels each)
    ("gaze_grcn", 256, 4, 4, 2, True, 1),      # 8-CTA clusters: 2 tiles x 4 channel quarters, training saves
    ("gaze_grcn", 128, 5, 2, 1, True, 1),      # 64-channel slices at hidden 128
    ("gaze_grcn", 256, 90, 4, 1, False, 1),    # 45 tiles x 4 = 180 CTAs on 148 SMs: clusters walk several tiles
])
def test_persistent_recurrence_matches_per_step_path(model, Ch, B, nsplit, mcast, train, ntile):
    """rgp_convgru_seq_fwd (all T steps in one launch) against the per-step tap-GEMM path on the same bf16 operands and
    gate functions.  The two accumulate in different orders (persistent kernel: tap-major; per-step slab GEMM: k-chunk
    major), so pre-activations differ in their last fp32 bits; where that flips the bf16 rounding of a state element the
    next step sees a one-ulp different operand.  States therefore agree to a fraction of a bf16 ulp, not to fp32 ulps."""
    T = 5
    p, c3d, _, gt = _case(model, min(B, 4), T, Ch)
    reps = (B + c3d.shape[0] - 1) // c3d.shape[0]
    c3d = np.concatenate([np.roll(c3d, i, axis=1) for i in range(reps)], axis=0)[:B]
    gt = np.concatenate([np.roll(gt, i, axis=1) for i in range(reps)], axis=0)[:B]
    out = []
    for seq in (False, True):
        eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=1, device="cuda", train=train, seq=seq, nsplit=nsplit, mcast=mcast,
                         ntile=ntile)
        assert eng.use_seq == seq and eng.ntile == ntile
        eng.load_tf_params(p)
        eng.forward(c3d=torch.from_numpy(c3d).cuda(), gt=torch.from_numpy(gt).cuda())
        torch.cuda.synchronize()
        d = dict(h=eng.Hf.cpu().numpy(), hop=eng.Hop.float().cpu().numpy(), logits=eng.logits.cpu().numpy())
        if train:
            d.update(z=eng.Z.cpu().numpy(), r=eng.Rs.cpu().numpy(), c=eng.Cs.cpu().numpy(), rh=eng.RH.float().cpu().numpy())
        out.append(d)
    a, b = out
    assert np.abs(a["h"]).max() > 1e-3
    for k in a:
        scale = np.abs(a[k]).max()
        # bf16 copies (hop, rh) may differ by one bf16 ulp where the fp32 values differ in their last bits
        tol = 2.0 ** -7 if k in ("hop", "rh") else (2e-3 if k == "logits" else 1e-3)
        assert np.abs(a[k] - b[k]).max() <= tol * scale + 1e-7, (k, np.abs(a[k] - b[k]).max(), scale)


@pytest.mark.parametrize("dtype", [0, 1])
def test_grid14_long_sequence_convlstm_states(dtype):
    """BASELINE config 4 family (ConvLSTM, 14x14 grid, long T): hidden and cell states after every step against the
    oracle's cell recurrence (gaze_lstm.py:103-133 as written).  14x14 has no reference head, so states are compared."""
    B, T, Ch, G, Cin = 2, 16, 128, 14, 512
    p = O.init_params("gaze_lstm", hidden=Ch, T=T, grid=G, seed=5, regime="S")
    rng = np.random.RandomState(11)
    emb = SY.make_embedded(rng, B, T, G, Cin)
    pe = dict(p)
    pe["out_W"], pe["out_b"] = np.zeros((Ch, 1)), np.zeros((1,))
    eng = GazeEngine("gaze_lstm", B=B, T=T, G=G, Ch=Ch, dtype=dtype, device="cuda", head="fc")
    eng.load_tf_params(pe)
    eng.forward(emb=torch.from_numpy(emb).cuda())
    torch.cuda.synchronize()
    pt = {k: torch.tensor(np.asarray(v), dtype=torch.float64) for k, v in p.items() if k.startswith("ConvLSTM")}
    state = torch.zeros(B, G, G, 2 * Ch, dtype=torch.float64)
    x = torch.from_numpy(emb).double()
    tol = 1e-4 if dtype == 0 else 2e-2
    for t in range(T):
        h, state = O.lstm_rcn_cell(x[:, t], state, pt)
        got = eng.read_state(t + 1)[1].cpu().double()
        assert (got - h).abs().max() <= tol * max(1.0, float(h.abs().max())), t
    c_fin = eng.read_state(T)[0].cpu().double()
    assert (c_fin - state[..., :Ch]).abs().max() <= tol * max(1.0, float(state.abs().max()))


def test_fc_head_backward_all_steps_matches_per_step_and_numpy():
    """rgp_fc_head_bwd_all (one launch for all T) against the per-step kernel and a float64 numpy statement of the same
    per-position FC backward: dhs[t, m] = dlogit * w, gW = sum dlogit * h, gb = sum dlogit."""
    lib = L.get_lib()
    B, T, G, Ch = 3, 5, 14, 128
    Wp, S = G + 1, (G + 1) ** 2
    M = B * S
    rng = np.random.RandomState(3)
    dl = torch.from_numpy(rng.randn(B, T, G, G).astype(np.float32)).cuda()
    h = torch.from_numpy(rng.randn((T + 1) * M + Wp + 2, Ch).astype(np.float32)).cuda()
    w = torch.from_numpy(rng.randn(Ch).astype(np.float32)).cuda()
    out = []
    for batched in (False, True):
        dhs = torch.zeros(T * M, Ch, device="cuda")
        gw, gb = torch.zeros(Ch, device="cuda"), torch.zeros(1, device="cuda")
        h1 = h.data_ptr() + M * Ch * 4
        if batched:
            lib.rgp_fc_head_bwd_all(dl.data_ptr(), h1, w.data_ptr(), dhs.data_ptr(), gw.data_ptr(), gb.data_ptr(), B, T, G, Ch, 0)
        else:
            for t in range(T):
                lib.rgp_fc_head_bwd(dl.data_ptr(), h1 + t * M * Ch * 4, w.data_ptr(), dhs.data_ptr() + t * M * Ch * 4,
                                    gw.data_ptr(), gb.data_ptr(), B, T, t, G, Ch, 0)
        torch.cuda.synchronize()
        out.append((dhs.cpu().numpy(), gw.cpu().numpy().astype(np.float64), gb.cpu().numpy().astype(np.float64)))
    assert np.array_equal(out[0][0], out[1][0])
    hn, dn, wn = h.cpu().numpy().astype(np.float64), dl.cpu().numpy().astype(np.float64), w.cpu().numpy().astype(np.float64)
    gw_ref, gb_ref = np.zeros(Ch), 0.0
    dhs_ref = np.zeros((T * M, Ch))
    for t in range(T):
        for b in range(B):
            for y in range(G):
                m = t * M + b * S + y * Wp + np.arange(G)
                d = dn[b, t, y]
                dhs_ref[m] = d[:, None] * wn[None]
                gw_ref += d @ hn[M + m + Wp + 1]
                gb_ref += d.sum()
    assert np.abs(out[1][0] - dhs_ref).max() <= 1e-6 * np.abs(dhs_ref).max()
    for gw_, gb_ in (out[0][1:], out[1][1:]):
        assert np.abs(gw_ - gw_ref).max() <= 1e-4 * np.abs(gw_ref).max() + 1e-4
        assert abs(gb_[0] - gb_ref) <= 1e-4 * abs(gb_ref) + 1e-3


@pytest.mark.parametrize("G,B,T,Ch,lstm", [(7, 3, 5, 128, False), (7, 37, 7, 256, False), (14, 2, 3, 128, True)])
def test_i2h_im2col_dense_rows_equal_slab_path(G, B, T, Ch, lstm, monkeypatch):
    """The input-half conv for all T frames in dense-M mode (TMA im2col loads, only the G*G real positions are GEMM rows)
    against the padded slab path: same pre-activations at every real position (fp32 accumulation order differs: 1e-5 of
    the tensor's scale), pad rows of Xg untouched (zero)."""
    import os
    model = "gaze_lstm" if lstm else "gaze_grcn77"
    p = O.init_params(model, hidden=Ch, T=T, grid=G, seed=4, regime="S", head="fc")
    emb = torch.from_numpy(SY.make_embedded(np.random.RandomState(3), B, T, G, 512)).cuda()
    out = []
    for flag in ("1", "0"):
        monkeypatch.setenv("RGP_IM2COL", flag)
        eng = GazeEngine(model, B=B, T=T, G=G, Ch=Ch, dtype=1, device="cuda", head="fc")
        eng.load_tf_params(p)
        eng.stage_embed(emb=emb)
        eng.stage_i2h()
        torch.cuda.synchronize()
        out.append(eng.Xg.view(T * B, eng.Wp, eng.Wp, -1).cpu().numpy())
    a, b = out
    scale = np.abs(b[:, :G, :G]).max()
    assert np.abs(a[:, :G, :G] - b[:, :G, :G]).max() <= 1e-5 * scale
    assert not a[:, G].any() and not a[:, :, G].any()          # dense mode never writes the pad row / column


@pytest.mark.parametrize("G,B,T,Ch,model", [(7, 40, 25, 128, "gaze_grcn"), (7, 48, 8, 256, "gaze_grcn"), (14, 16, 16, 128, "gaze_lstm")])
def test_position_major_conv_tiles_equal_dense_rows_bitwise(G, B, T, Ch, model, monkeypatch):
    """Position-major tiles of the 3x3 SAME conv (im2col 3 / 4: one position of 128 frames per tile, out-of-frame taps
    skipped) against the dense-M order (im2col 1 / 2): the skipped taps only ever added zeros and the order of the
    remaining ones is the same, so the i2h pre-activations and the BPTT's dX of the input half are equal BIT FOR BIT -
    ragged last frame group (1000 = 7 * 128 + 104 frames) and the 14x14 grid included."""
    head = None if G == 7 else "fc"
    p = O.init_params(model, hidden=Ch, T=T, grid=G, seed=4, regime="S", head=head)
    rng = np.random.RandomState(3)
    emb = torch.from_numpy(SY.make_embedded(rng, B, T, G, 512)).cuda()
    GH = G if head == "fc" else 49
    gt = torch.from_numpy(SY.make_gazemaps(rng, B, T, GH, GH, normalize=True)[0]).cuda()
    out = []
    for flag in ("1", "0"):
        monkeypatch.setenv("RGP_PMAJOR", flag)
        eng = GazeEngine(model, B=B, T=T, G=G, Ch=Ch, dtype=1, device="cuda", head=head, train=True)
        eng.load_tf_params(p)
        eng.forward(emb=emb, gt=gt)
        xg = eng.Xg.clone()
        lg = eng.logits.clone()      # 7x7 models: the composed head GEMM takes position-major tiles too (im2col 5)
        # the transposed conv (dX of the input half) on the step's dG, called as the c3d entry's backward does
        btaps = [eng.Wp + 1 - o for o in eng.taps3]
        eng.backward()
        NG = eng.ngx * Ch
        eng.lib.rgp_transpose_taps(eng._ptr("Wx_op"), eng.WxT.data_ptr(), 9, NG, eng.Cin, eng.dtype, eng._s())
        dx = torch.zeros(T * B * G * G, eng.Cin, dtype=torch.bfloat16, device="cuda")
        eng._gemm(A=eng.dG.data_ptr(), rowsA=T * B * eng.S, lda=NG, K=NG, Bw=eng.WxT.data_ptr(), M=T * B * G * G, N=eng.Cin,
                  taps=btaps, epi=L.EPI_DX_DENSE, S=eng.S, Wp=eng.Wp, G=G, P=G * G, out_op=dx.data_ptr(), ldo=eng.Cin, im2col=2)
        torch.cuda.synchronize()
        out.append((xg.cpu().numpy(), dx.float().cpu().numpy(), lg.cpu().numpy()))
    (xa, da, la), (xb, db, lb) = out
    assert np.abs(xb).max() > 0 and np.abs(db).max() > 0
    assert np.array_equal(xa, xb)
    assert np.array_equal(da, db)
    assert np.abs(lb).max() > 0 and np.array_equal(la, lb)
