"""GPU tests of the optimiser kernels (base.py:262-297, gaze_rnn.py:435-444) and of the graph-captured training step.

* rgp_sumsq + rgp_gradnorm_clip_adam on known (g, p, m, v) against the oracle's clip_by_global_norm + adam_step, to
  fp32 rounding, on all three clip branches (||g|| above the clip, below it, clipping disabled) at t = 1 and t = 1000;
* rgp_adam_schedule (device-side global_step, staircase learning rate, Adam bias correction) against the host formula;
* a CUDA graph of fwd + BPTT + clip + Adam replayed 10 times == 10 eager optimiser steps (same schedule, same step
  counter, parameters equal up to the summation order of the split-K weight-gradient atomics);
* bf16-fed features give bit-identical results to fp32-fed features on the bf16 engine.
"""
import math

import numpy as np
import pytest
import torch

from oracle import model_oracle as O
from recurrent_gaze_prediction_b200 import lib as L
from recurrent_gaze_prediction_b200 import synthetic as SY
from recurrent_gaze_prediction_b200.engine import GazeEngine, StepGraph

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("clip,gscale", [(10.0, 100.0), (10.0, 1e-3), (0.0, 100.0), (-1.0, 1.0)])
@pytest.mark.parametrize("t", [1, 1000])
def test_clip_adam_kernel_matches_oracle(clip, gscale, t):
    lib = L.get_lib()
    rng = np.random.RandomState(int(t + gscale))
    n = 100003
    g = (rng.standard_normal(n) * gscale / math.sqrt(n)).astype(np.float32)
    p = rng.standard_normal(n).astype(np.float32)
    m = (rng.standard_normal(n) * 0.01).astype(np.float32)
    v = (rng.uniform(size=n) * 1e-4).astype(np.float32)
    lr = 1e-3
    lr_t = lr * math.sqrt(1.0 - 0.999 ** t) / (1.0 - 0.9 ** t)
    if clip > 0:
        (gc,), norm = O.clip_by_global_norm([g.astype(np.float64)], clip)
        assert (norm > clip) == (gscale > 1.0)          # both branches of g*clip/max(||g||, clip) are exercised
    else:
        gc = g.astype(np.float64)
    # TF casts the hyper-parameters to the variable dtype before ApplyAdam ([TF-lib] adam.py: math_ops.cast(beta2_t,
    # var.dtype)), so 1 - beta2 is formed from float32(0.999) = 0.99900001287: the oracle is given those values
    # (lr_t itself is an input of the kernel and is passed identically to both sides)
    b1, b2 = float(np.float32(0.9)), float(np.float32(0.999))
    (p_ref,), (m_ref,), (v_ref,) = O.adam_step([p.astype(np.float64)], [gc], [m.astype(np.float64)],
                                               [v.astype(np.float64)], t, lr, beta1=b1, beta2=b2)
    lr_nom = lr * math.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
    p_ref = p.astype(np.float64) + (p_ref - p.astype(np.float64)) * (lr_t / lr_nom)
    d = [torch.from_numpy(x).cuda() for x in (p, g, m, v)]
    sumsq = torch.zeros(1, dtype=torch.float64, device="cuda")
    p_op = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    lib.rgp_sumsq(d[1].data_ptr(), n, sumsq.data_ptr(), s)
    lib.rgp_gradnorm_clip_adam(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), n, sumsq.data_ptr(),
                               clip, lr_t, 0.9, 0.999, 1e-8, p_op.data_ptr(), 1, s)
    torch.cuda.synchronize()
    assert abs(float(sumsq.cpu()[0]) - float((g.astype(np.float64) ** 2).sum())) <= 1e-10 * float((g.astype(np.float64) ** 2).sum())
    pg, mg, vg = (x.cpu().numpy().astype(np.float64) for x in (d[0], d[2], d[3]))
    assert np.abs(mg - m_ref).max() <= 4e-7 * max(1e-30, np.abs(m_ref).max())
    assert np.abs(vg - v_ref).max() <= 4e-7 * max(1e-30, np.abs(v_ref).max())
    # the update itself: a few fp32 ulps of the step size lr_t * m / (sqrt(v) + eps), plus one ulp of |p|
    upd = np.abs(p.astype(np.float64) - p_ref)
    assert (np.abs(pg - p_ref) <= 1e-5 * upd + 1.2e-7 * np.abs(p_ref) + 1e-12).all()
    assert torch.equal(p_op.cpu(), d[0].cpu().to(torch.bfloat16))       # the operand copy is the rounded new parameter


def test_device_schedule_matches_host_formula():
    lib = L.get_lib()
    step = torch.zeros(1, dtype=torch.int64, device="cuda")
    lr0 = torch.full((1,), 3e-3, dtype=torch.float64, device="cuda")
    sched = torch.zeros(2, dtype=torch.float32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for start in (0, 497, 123456):
        step.fill_(start)
        for k in range(6):
            lib.rgp_adam_schedule(step.data_ptr(), lr0.data_ptr(), 0.8, 500, 0.9, 0.999, sched.data_ptr(), s)
            st = start + k
            lr = O.learning_rate(st, 3e-3, 0.8, 500)
            lr_t = lr * math.sqrt(1.0 - 0.999 ** (st + 1)) / (1.0 - 0.9 ** (st + 1))
            got = sched.cpu().numpy()
            assert int(step.cpu()[0]) == st + 1
            assert abs(got[1] - lr) <= 6e-8 * lr and abs(got[0] - lr_t) <= 1.2e-7 * lr_t, (st, got, lr_t, lr)


@pytest.mark.parametrize("model,Ch", [("gaze_grcn", 128), ("gaze_lstm", 128)])
def test_graph_replays_equal_eager_optimizer_steps(model, Ch):
    """10 replays of ONE captured graph (fwd + BPTT + clip + Adam) against 10 eager steps: identical device-side step
    counter and (lr_t, lr) after every step (bit-equal: same kernel, same inputs), decayed learning rate included
    (decay_steps = 4 so the staircase moves inside the 10 steps); parameters agree to the summation-order noise of the
    weight-gradient atomics."""
    B, T = 4, 5
    p = O.init_params(model, hidden=Ch, T=T, seed=0, regime="S")
    rng = np.random.RandomState(3)
    c3d = torch.from_numpy(SY.make_c3d(rng, B, T)).cuda()
    gt = torch.from_numpy(SY.make_gazemaps(rng, B, T, 49, 49, normalize=True)[0]).cuda()
    opt = dict(lr0=1e-3, decay=0.5, clip=10.0, decay_steps=4)
    engs, scheds = [], []
    for graph in (False, True):
        eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=1, device="cuda", train=True)
        eng.load_tf_params(p)
        fwd = lambda e=eng: e.forward(c3d=c3d, gt=gt)
        sg = StepGraph(eng, fwd, train=True, opt=opt) if graph else None
        if graph:
            assert eng.step == 0 and eng.step_on_device() == 0        # building the graph did not train
            assert torch.equal(eng.params.cpu(), engs[0][1])
        else:
            engs.append((eng, eng.params.cpu().clone()))
        seq = []
        for k in range(10):
            if graph:
                sg.replay()
            else:
                fwd()
                eng.backward()
                eng.optimizer_step(**opt)
            seq.append(eng.sched.cpu().numpy().copy())
        torch.cuda.synchronize()
        assert eng.step == 10 and eng.step_on_device() == 10
        scheds.append(np.stack(seq))
        if graph:
            engs.append((eng, None))
    assert np.array_equal(scheds[0], scheds[1])
    assert scheds[0][3, 1] == np.float32(1e-3) and scheds[0][4, 1] == np.float32(5e-4) and scheds[0][8, 1] == np.float32(2.5e-4)
    a, b = engs[0][0].params.cpu().numpy(), engs[1][0].params.cpu().numpy()
    p0 = engs[0][1].numpy()
    assert np.abs(a - p0).max() > 1e-3                                   # ten real updates happened
    # Adam divides by sqrt(v): where the true gradient is ~0 (out_b: sum(softmax - gt) == 0) the atomics' summation-order
    # noise IS the gradient and its sign decides a +-lr step, so the bar is on the update as a whole (L2) and loosely
    # on the worst element
    upd = (a - p0).astype(np.float64)
    assert np.linalg.norm((a - b).astype(np.float64)) <= 2e-3 * np.linalg.norm(upd)
    assert np.abs(a - b).max() <= 5e-2 * np.abs(upd).max()


def test_bf16_fed_features_are_bit_identical_to_fp32_fed():
    B, T, Ch = 3, 4, 128
    p = O.init_params("gaze_grcn", hidden=Ch, T=T, seed=1, regime="S")
    rng = np.random.RandomState(5)
    c3d = torch.from_numpy(SY.make_c3d(rng, B, T)).cuda()
    out = []
    for src in (c3d, c3d.to(torch.bfloat16)):
        eng = GazeEngine("gaze_grcn", B=B, T=T, Ch=Ch, dtype=1, device="cuda")
        eng.load_tf_params(p)
        eng.forward(c3d=src)
        torch.cuda.synchronize()
        out.append((eng.Apack.clone(), eng.logits.clone()))
    assert torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1])
    # fp32 engine: bf16-fed features are the bf16-rounded features
    eng = GazeEngine("gaze_grcn", B=B, T=T, Ch=Ch, dtype=0, device="cuda")
    eng.load_tf_params(p)
    eng.stage_embed(c3d.to(torch.bfloat16))
    a = eng.Apack.clone()
    eng.stage_embed(c3d.to(torch.bfloat16).float())
    assert torch.equal(a, eng.Apack)


@pytest.mark.parametrize("kind", ["sgd", "rmsprop"])
@pytest.mark.parametrize("clip,gscale", [(10.0, 100.0), (10.0, 1e-3), (0.0, 100.0)])
def test_clip_momentum_rmsprop_kernel_matches_oracle(kind, clip, gscale):
    """rgp_gradnorm_clip_momentum_dev (models/base.py:270-273: RMSProp(momentum 0.9) / Momentum(0.9) in TF1's update rules)
    on known (g, p, slots) against the oracle's clip_by_global_norm + momentum_step / rmsprop_step, three consecutive steps,
    to float32 rounding; the bf16 operand copy equals the rounded parameters."""
    lib = L.get_lib()
    rng = np.random.RandomState(7)
    n = 70001
    p = rng.standard_normal(n).astype(np.float32)
    acc = np.zeros(n, np.float32)
    ms = np.ones(n, np.float32)
    d = [torch.from_numpy(x.copy()).cuda() for x in (p, acc, ms)]
    pr, ar, mr = p.astype(np.float64), acc.astype(np.float64), ms.astype(np.float64)
    sched = torch.tensor([0.0, 1e-3], dtype=torch.float32, device="cuda")
    lr = float(np.float32(1e-3))
    p_op = torch.zeros(n, dtype=torch.bfloat16, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    mom, dec = float(np.float32(0.9)), float(np.float32(0.9))
    for step in range(3):
        g = (rng.standard_normal(n) * gscale / math.sqrt(n)).astype(np.float32)
        gd = torch.from_numpy(g).cuda()
        sumsq = torch.zeros(1, dtype=torch.float64, device="cuda")
        lib.rgp_sumsq(gd.data_ptr(), n, sumsq.data_ptr(), s)
        lib.rgp_gradnorm_clip_momentum_dev(d[0].data_ptr(), gd.data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n,
                                           sumsq.data_ptr(), clip, sched.data_ptr(), 1 if kind == "rmsprop" else 0, 0.9, 0.9,
                                           1e-10, p_op.data_ptr(), 1, s)
        gc = O.clip_by_global_norm([g.astype(np.float64)], clip)[0][0] if clip > 0 else g.astype(np.float64)
        p_before = pr.copy()
        if kind == "rmsprop":
            (pr,), (mr,), (ar,) = O.rmsprop_step([pr], [gc], [mr], [ar], lr, decay=dec, momentum=mom)
        else:
            (pr,), (ar,) = O.momentum_step([pr], [gc], [ar], lr, momentum=mom)
        torch.cuda.synchronize()
        pg, ag = d[0].cpu().numpy().astype(np.float64), d[1].cpu().numpy().astype(np.float64)
        assert np.abs(ag - ar).max() <= 1e-6 * max(np.abs(ar).max(), 1e-30), (step, np.abs(ag - ar).max())
        assert np.abs(pg - pr).max() <= 2e-7 * np.abs(pr).max() + 1e-6 * np.abs(pr - p_before).max()
        if kind == "rmsprop":
            mg = d[2].cpu().numpy().astype(np.float64)
            assert np.abs(mg - mr).max() <= 4e-7 * np.abs(mr).max()
        assert torch.equal(p_op, d[0].to(torch.bfloat16))
        pr, ar, mr = pg, ag, (d[2].cpu().numpy().astype(np.float64) if kind == "rmsprop" else mr)   # no drift between steps
