#!/usr/bin/env python
"""Small shapes of the hand-rolled mbarrier / TMA / TMEM kernels, to be run under compute-sanitizer:

    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python tools/sanitize_cases.py [case ...]

Cases: tapgemm256 (tapgemm_tc_kernel<256,true>, slab mode, >148 tiles), tapgemm128 (per-step recurrence shape),
head48 (tapgemm_tc_kernel<48,false>), wgrad (wgrad_tc_kernel<256>), seq_<nsplit>_<mcast> (gru_seq_kernel, cluster
shapes 1x1, 2x1, 1x4, 2x4), bptt (one small fwd + BPTT + Adam step).  Each prints 'case <name> ok' after a device sync."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from recurrent_gaze_prediction_b200 import lib as L
from recurrent_gaze_prediction_b200.engine import GazeEngine
from recurrent_gaze_prediction_b200.models.model_util import init_parameters
from recurrent_gaze_prediction_b200 import synthetic as SY


def tapgemm(M, N, K, taps, epi=L.EPI_STORE_F32):
    rowsA = M + max(taps) + 8
    A = torch.randn(rowsA, K, device="cuda").bfloat16()
    Bw = torch.randn(len(taps) * N, K, device="cuda").bfloat16()
    out = torch.zeros(M, N, device="cuda")
    a = L.TapGemmArgs()
    a.A, a.rowsA, a.lda, a.Bw = A.data_ptr(), rowsA, K, Bw.data_ptr()
    a.M, a.N, a.K, a.ntaps = M, N, K, len(taps)
    for i, o in enumerate(taps):
        a.tap_off[i] = o
    a.dtype, a.engine, a.epi, a.out_f32, a.ldo = 1, 0, epi, out.data_ptr(), N
    L.get_lib().rgp_tapgemm(C.byref(a), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()


def wgrad():
    M, Ny, Nx, taps = 1000, 256, 256, [0, 1, 2, 8, 9, 10, 16, 17, 18]
    Y = torch.randn(M, Ny, device="cuda").bfloat16()
    X = torch.randn(M + 32, Nx, device="cuda").bfloat16()
    D = torch.zeros(len(taps), Ny, Nx, device="cuda")
    a = L.WgradArgs()
    a.Y, a.ldy, a.ycol0, a.Ny = Y.data_ptr(), Ny, 0, Ny
    a.X, a.rowsX, a.ldx, a.xcol0, a.Nx = X.data_ptr(), M + 32, Nx, 0, Nx
    a.M, a.ntaps = M, len(taps)
    for i, o in enumerate(taps):
        a.tap_off[i] = o
    a.D, a.dtype, a.engine = D.data_ptr(), 1, 0
    L.get_lib().rgp_wgrad(C.byref(a), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()


def engine_case(seq, nsplit=1, mcast=1, train=False, B=5, T=2, Ch=256, model="gaze_grcn"):
    p = init_parameters(model, hidden=Ch, T=T, seed=0)
    rng = np.random.RandomState(0)
    eng = GazeEngine(model, B=B, T=T, Ch=Ch, dtype=1, device="cuda", train=train, seq=seq, nsplit=nsplit, mcast=mcast)
    eng.load_tf_params(p)
    c3d = torch.from_numpy(SY.make_c3d(rng, B, T)).cuda()
    gt = torch.from_numpy(SY.make_gazemaps(rng, B, T, 49, 49, normalize=True)[0]).cuda()
    eng.forward(c3d=c3d, gt=gt)
    if train:
        eng.backward()
        eng.optimizer_step()
    torch.cuda.synchronize()


TAPS9 = [0, 1, 2, 8, 9, 10, 16, 17, 18]
CASES = {
    "tapgemm256": lambda: tapgemm(150 * 128, 256, 64, TAPS9),
    "tapgemm128": lambda: tapgemm(1000, 512, 128, TAPS9),
    "head48": lambda: tapgemm(515, 48, 128, [33 - 10 * dy - dx for dy in range(4) for dx in range(4)]),
    "wgrad": wgrad,
    "seq_1_1": lambda: engine_case(True, 1, 1),
    "seq_2_1": lambda: engine_case(True, 2, 1),
    "seq_1_4": lambda: engine_case(True, 1, 4, B=14),
    "seq_2_4": lambda: engine_case(True, 2, 4, B=14, train=True),
    "bptt": lambda: engine_case(False, train=True, B=3, Ch=128),
    "lstm": lambda: engine_case(False, train=True, B=3, Ch=128, model="gaze_lstm"),
}

if __name__ == "__main__":
    torch.manual_seed(0)
    for name in (sys.argv[1:] or list(CASES)):
        CASES[name]()
        print("case %s ok" % name, flush=True)
