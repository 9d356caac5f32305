"""DenseGruEngine — the dense-GRU variant of the gaze model (reference ``models/gaze_rnn.py:211-361``)
over the C-ABI: 1x1 projection 1024 -> 32 per position, the [7,7,32] map flattened (y, x, c) into one dense GRU of
7*7*32 + 49 = 1617 units (``:243-246,315``), ``proj_out`` [1617, GH*GW] per step (``:318-319,346-348``).

Every contraction is a plain GEMM on the tap-GEMM engines (``rgp_tapgemm`` with one tap): the projection and the
input halves of all three gate blocks for all T frames at once, two small GEMMs per step for the recurrent halves, and
the output projection for all T at once.  Widths are padded to the engine's granularity (I 1568 -> 1600, H 1617 -> 1664,
GH*GW 2401 -> 2432) with zero weights; the padded state columns stay exactly zero.  The element-wise glue is
``csrc/dense_gru.cu``.

Training (``train=True``; the reference trains this variant through ``tf.gradients`` over ``tf.nn.rnn_cell.GRUCell``,
``gaze_rnn.py:447-478`` + ``base.py:262-297``): hand-written BPTT on the same engines - the output projection's dX / dW for
all T at once, per step (reverse time) two element-wise gate-backward kernels around two transposed-weight GEMMs, then
the cell's weight gradients as three row contractions over all T*B rows (``rgp_wgrad``), dX of the input halves, the
projection's gradients; clip + TF-Adam through :class:`TrainStateMixin`.  Dropout on the embedded features
(``:302-303``, keep_prob 0.5 while training ``:529``) is a counter-based mask re-derived in the backward pass.

The ShallowNet frame-saliency branch the reference also builds (``:252-276``) never reaches the prediction — its entry in
the RNN input concat is commented out (``:333``) — so ``frame_images`` is not consumed.
"""
from __future__ import annotations

import ctypes as C
from collections import OrderedDict

import numpy as np
import torch

from . import lib as L
from .engine import LOSS_TYPES, TrainStateMixin, _tdtype


def _pad(n, m):
    return (n + m - 1) // m * m


class DenseGruEngine(TrainStateMixin):
    model = "gaze_rnn"
    lstm = False
    use_bn = False
    use_seq = False

    def __init__(self, B=2, T=35, G=7, Cp=32, Cfeat=1024, GH=49, dtype=0, device="cuda", lib=None, train=False,
                 engine=None):
        self.B, self.T, self.G, self.Cp, self.Cfeat, self.GH = B, T, G, Cp, Cfeat, GH
        self.P = G * G
        self.I = self.P * Cp                       # gaze_rnn.py:244
        self.H = self.I + self.P                   # gaze_rnn.py:245: "C3D projected PLUS saliency map (49)"
        self.Ch = self.H
        self.N = GH * GH
        self.Ip, self.Hp, self.Np = _pad(self.I, 64), _pad(self.H, 128), _pad(self.N, 128)
        self.dtype, self.train, self.engine = dtype, bool(train), engine or 0
        self.keep_prob, self.dropout_seed = 1.0, 0          # set per step by the model class while training
        self.param_version = 0
        self.device = torch.device(device)
        self.lib = lib if lib is not None else L.get_lib()
        self.tdt = _tdtype(dtype)
        self._build_params()
        if self.train:
            self._alloc_train_state()
        self._alloc()

    # ------------------------------------------------------------------ parameters (flat fp32 + operand copy)
    def _build_params(self):
        spec = OrderedDict()
        spec["proj_W_op"] = (self.Cp, self.Cfeat)            # [N, K] K contiguous
        spec["proj_b"] = (self.Cp,)
        spec["Wx_op"] = (3 * self.Hp, self.Ip)               # rows r | u | c, x-halves of the two GRUCell kernels
        spec["Whg_op"] = (2 * self.Hp, self.Hp)              # rows r | u, h-half of gates/kernel
        spec["Whc_op"] = (self.Hp, self.Hp)                  # h-half of candidate/kernel
        spec["bias_g"] = (2 * self.Hp,)
        spec["bias_c"] = (self.Hp,)
        spec["out_W_op"] = (self.Np, self.Hp)
        spec["out_b"] = (self.Np,)
        self.spec, self.offsets = spec, OrderedDict()
        off = 0
        for k, shp in spec.items():
            n = int(np.prod(shp))
            self.offsets[k] = (off, n)
            off += _pad(n, 64)
        self.nparams = off
        self.params = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.params_op = self.params if self.dtype == 0 else torch.zeros(off, dtype=self.tdt, device=self.device)

    def view(self, name, buf=None):
        off, n = self.offsets[name]
        buf = self.params if buf is None else buf
        return buf[off:off + n].view(*self.spec[name])

    def _ptr(self, name, buf=None):
        off, _ = self.offsets[name]
        buf = self.params if buf is None else buf
        return buf.data_ptr() + off * buf.element_size()

    def _pop(self, name):
        return self._ptr(name, self.params_op)

    def load_tf_params(self, p):
        """TF variables (reference names / shapes) -> padded op layouts.  gates/kernel [(I+H), 2H] columns r | u."""
        I, H, Ip, Hp, N, Np = self.I, self.H, self.Ip, self.Hp, self.N, self.Np
        f = lambda k: np.asarray(p[k], dtype=np.float32)
        host = {k: np.zeros(s, dtype=np.float32) for k, s in self.spec.items()}
        host["proj_W_op"][:] = f("proj_c3d_W").T
        host["proj_b"][:] = f("proj_c3d_b")
        Wg, Wc = f("RNN/gru_cell/gates/kernel"), f("RNN/gru_cell/candidate/kernel")
        assert Wg.shape == (I + H, 2 * H) and Wc.shape == (I + H, H), (Wg.shape, Wc.shape)
        for blk, W in ((0, Wg[:, :H]), (1, Wg[:, H:]), (2, Wc)):
            host["Wx_op"][blk * Hp:blk * Hp + H, :I] = W[:I].T
        host["Whg_op"][:H, :H] = Wg[I:, :H].T
        host["Whg_op"][Hp:Hp + H, :H] = Wg[I:, H:].T
        host["Whc_op"][:H, :H] = Wc[I:].T
        bg = f("RNN/gru_cell/gates/bias")
        host["bias_g"][:H], host["bias_g"][Hp:Hp + H] = bg[:H], bg[H:]
        host["bias_c"][:H] = f("RNN/gru_cell/candidate/bias")
        host["out_W_op"][:N, :H] = f("proj_out_W").T
        host["out_b"][:N] = f("proj_out_b")
        flat = np.zeros(self.nparams, dtype=np.float32)
        for k, v in host.items():
            off, n = self.offsets[k]
            flat[off:off + n] = v.reshape(-1)
        self.params.copy_(torch.from_numpy(flat))
        self.refresh_operands()

    def pad_mask(self):
        """1 where a flat-buffer element is a real parameter, 0 in the zero padding (padded rows / columns of the op
        layouts).  The padding's gradients are exactly zero, so Adam never moves it; used by tests."""
        m = {k: np.zeros(s, dtype=np.float32) for k, s in self.spec.items()}
        I, H, Hp, N = self.I, self.H, self.Hp, self.N
        m["proj_W_op"][:] = 1
        m["proj_b"][:] = 1
        for blk in range(3):
            m["Wx_op"][blk * Hp:blk * Hp + H, :I] = 1
        m["Whg_op"][:H, :H] = 1
        m["Whg_op"][Hp:Hp + H, :H] = 1
        m["Whc_op"][:H, :H] = 1
        m["bias_g"][:H] = 1
        m["bias_g"][Hp:Hp + H] = 1
        m["bias_c"][:H] = 1
        m["out_W_op"][:N, :H] = 1
        m["out_b"][:N] = 1
        flat = np.zeros(self.nparams, dtype=np.float32)
        for k, v in m.items():
            off, n = self.offsets[k]
            flat[off:off + n] = v.reshape(-1)
        return flat

    def op_to_tf(self, flat):
        I, H, Hp, N = self.I, self.H, self.Hp, self.N
        flat = np.asarray(flat)
        g = lambda k: flat[self.offsets[k][0]:self.offsets[k][0] + self.offsets[k][1]].reshape(self.spec[k])
        out = OrderedDict()
        out["proj_c3d_W"] = g("proj_W_op").T.copy()
        out["proj_c3d_b"] = g("proj_b").copy()
        Wx, Whg, Whc = g("Wx_op"), g("Whg_op"), g("Whc_op")
        out["RNN/gru_cell/gates/kernel"] = np.concatenate(
            [np.concatenate([Wx[:H, :I].T, Wx[Hp:Hp + H, :I].T], axis=1),
             np.concatenate([Whg[:H, :H].T, Whg[Hp:Hp + H, :H].T], axis=1)], axis=0)
        out["RNN/gru_cell/gates/bias"] = np.concatenate([g("bias_g")[:H], g("bias_g")[Hp:Hp + H]])
        out["RNN/gru_cell/candidate/kernel"] = np.concatenate([Wx[2 * Hp:2 * Hp + H, :I].T, Whc[:H, :H].T], axis=0)
        out["RNN/gru_cell/candidate/bias"] = g("bias_c")[:H].copy()
        out["proj_out_W"] = g("out_W_op")[:N, :H].T.copy()
        out["proj_out_b"] = g("out_b")[:N].copy()
        return out

    def refresh_operands(self):
        if self.dtype != 0:
            self.lib.rgp_cast(self.params.data_ptr(), self.params_op.data_ptr(), self.nparams, self.dtype, self._s())
        self.param_version += 1

    # ------------------------------------------------------------------ buffers
    def _alloc(self):
        B, T, dev = self.B, self.T, self.device
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)
        rows = _pad(T * B, 128) + 128                         # whole 128-row TMA boxes stay inside the allocation
        self.Apack = z(T * B * self.P + 128, self.Cfeat, dt=self.tdt)
        self.Pacc = z(T * B * self.P, self.Cp)
        self.X = z(rows, self.Ip, dt=self.tdt)
        self.Xg = z(T * B, 3 * self.Hp)
        # h_0..h_T: fp32 master and operand copy (one buffer when the operands are fp32); +256 rows so that the 128-row
        # box of the last step's tile stays inside the allocation
        if self.dtype == 0:
            self.Hf = self.Hop = z((T + 1) * B + 256, self.Hp)
        else:
            self.Hf = z((T + 1) * B, self.Hp)
            self.Hop = z((T + 1) * B + 256, self.Hp, dt=self.tdt)
        # r*h: one slab (inference) or one per step (training: the X operand of dW_candidate)
        self.RH = z((T * B if self.train else 0) + _pad(B, 128) + 128, self.Hp, dt=self.tdt)
        self.Hg = z(B, 2 * self.Hp)
        self.Hc = z(B, self.Hp)
        self.U = z(B, self.Hp)
        self.Oacc = z(T * B, self.Np)
        self.logits = z(B, T, self.GH, self.GH)
        self.probs = z(B, T, self.GH, self.GH)
        self.frame_loss = z(B * T)
        self.loss = z(1)
        if self.train:
            Hp, Ip, Np = self.Hp, self.Ip, self.Np
            self.Us, self.Rs, self.Cs = z(T * B, Hp), z(T * B, Hp), z(T * B, Hp)      # saved gates of every step
            self.dlogits = z(B, T, self.GH, self.GH)
            self.dL = z(rows, Np, dt=self.tdt)                 # dlogits as time-major operand rows
            self.dHhead = z(T * B, Hp)                         # dL . W_out^T: head term of dL/dh_{t+1}
            self.dG = z(rows, 3 * Hp, dt=self.tdt)             # d(pre-activations) r | u | c of every step
            self.dhA, self.dhB = z(B, Hp), z(B, Hp)            # dL/dh (recurrent part), ping-pong
            self.drh = z(B, Hp)
            self.dX = z(T * B, Ip)
            self.dPacc = z(T * B * self.P + 128, self.Cp, dt=self.tdt)
            self.outWT = z(Hp, Np, dt=self.tdt)                # transposed operand copies (dX = dY . W)
            self.WhcT = z(Hp, Hp, dt=self.tdt)
            self.WhgT = z(Hp, 2 * Hp, dt=self.tdt)
            self.WxT = z(Ip, 3 * Hp, dt=self.tdt)

    def _s(self):
        if self.device.type == "cuda":
            return torch.cuda.current_stream(self.device).cuda_stream
        return 0

    def _gemm(self, A, rowsA, lda, Bw, M, N, K, out, ldo, accumulate=0):
        a = L.TapGemmArgs()
        a.dtype, a.engine = self.dtype, self.engine
        a.ntaps = 1
        a.tap_off[0] = 0
        a.A, a.rowsA, a.lda, a.Bw, a.M, a.N, a.K = A, rowsA, lda, Bw, M, N, K
        a.epi, a.out_f32, a.ldo, a.accumulate = L.EPI_STORE_F32, out, ldo, accumulate
        a.S, a.Wp, a.G = 1, 1, 1
        self.lib.rgp_tapgemm(C.byref(a), self._s())

    def reset_state(self):
        pass

    # ------------------------------------------------------------------ forward
    def forward(self, c3d=None, emb=None, gt=None, loss_type="xentropy", global_batch=None, carry_state=False, t0=0):
        """c3d [B,T,1024,7,7] fp32 -> logits [B,T,GH,GH]; with gt also probs and the loss (gaze_rnn.py:149-159,363-408)."""
        assert emb is None and not carry_state and t0 == 0, "the dense variant takes C3D features from h_0 = 0"
        lib, s = self.lib, self._s()
        B, T, P, Cp, Hp, Ip, H = self.B, self.T, self.P, self.Cp, self.Hp, self.Ip, self.H
        assert tuple(c3d.shape) == (B, T, self.Cfeat, self.G, self.G) and c3d.dtype == torch.float32
        es = self.Hop.element_size()
        # (1) layout + 1x1 projection (gaze_rnn.py:285-311), then the flattened GRU input rows (:335-338)
        lib.rgp_pack_c3d(c3d.data_ptr(), self.Apack.data_ptr(), B, T, self.Cfeat, P, self.dtype, s)
        self._gemm(self.Apack.data_ptr(), self.Apack.shape[0], self.Cfeat, self._pop("proj_W_op"), T * B * P, Cp,
                   self.Cfeat, self.Pacc.data_ptr(), Cp)
        lib.rgp_dense_pack_dropout(self.Pacc.data_ptr(), Cp, self._ptr("proj_b"), self.X.data_ptr(), Ip, T * B, P, Cp,
                                   float(self.keep_prob), int(self.dropout_seed), self.dtype, s)
        # (2) input halves of r | u | c for all T frames at once
        self._gemm(self.X.data_ptr(), self.X.shape[0], Ip, self._pop("Wx_op"), T * B, 3 * Hp, Ip, self.Xg.data_ptr(),
                   3 * Hp)
        # (3) the recurrence (gaze_rnn.py:324-341), h_0 = 0
        lib.rgp_zero(self.Hf.data_ptr(), B * Hp * 4, s)
        if self.Hop is not self.Hf:
            lib.rgp_zero(self.Hop.data_ptr(), B * Hp * es, s)
        rowsH = self.Hop.shape[0]
        for t in range(T):
            hf = self.Hf.data_ptr() + t * B * Hp * 4
            hop = self.Hop.data_ptr() + t * B * Hp * es
            xg = self.Xg.data_ptr() + t * B * 3 * Hp * 4
            sl = t * B * Hp * 4
            u = (self.Us.data_ptr() + sl) if self.train else self.U.data_ptr()
            rh = self.RH.data_ptr() + (t * B * Hp * es if self.train else 0)
            rh_rows = self.RH.shape[0] - (t * B if self.train else 0)
            self._gemm(hop, rowsH - t * B, Hp, self._pop("Whg_op"), B, 2 * Hp, Hp, self.Hg.data_ptr(), 2 * Hp)
            lib.rgp_dense_gru_ru(xg, 3 * Hp, self.Hg.data_ptr(), self._ptr("bias_g"), hf, rh, u,
                                 (self.Rs.data_ptr() + sl) if self.train else 0, B, H, Hp, self.dtype, s)
            self._gemm(rh, rh_rows, Hp, self._pop("Whc_op"), B, Hp, Hp, self.Hc.data_ptr(), Hp)
            lib.rgp_dense_gru_c(xg, 3 * Hp, self.Hc.data_ptr(), self._ptr("bias_c"), u, hf,
                                hf + B * Hp * 4, hop + B * Hp * es, (self.Cs.data_ptr() + sl) if self.train else 0,
                                B, H, Hp, self.dtype, s)
        # (4) proj_out for all T at once (gaze_rnn.py:346-359)
        self._gemm(self.Hop.data_ptr() + B * Hp * es, rowsH - B, Hp, self._pop("out_W_op"), T * B, self.Np, Hp,
                   self.Oacc.data_ptr(), self.Np)
        lib.rgp_dense_logits(self.Oacc.data_ptr(), self.Np, self._ptr("out_b"), self.logits.data_ptr(), B, T, self.N, s)
        self.entry = "c3d"
        if gt is not None:
            self.loss_fwd(gt, loss_type, global_batch)
        return self.logits

    def loss_fwd(self, gt, loss_type="xentropy", global_batch=None):
        if loss_type not in LOSS_TYPES:
            raise NotImplementedError(str(loss_type))
        assert tuple(gt.shape) == tuple(self.logits.shape) and gt.dtype == torch.float32
        n, hw = self.B * self.T, self.N
        self.lib.rgp_softmax_xent_fwd_bwd(self.logits.data_ptr(), gt.data_ptr(), self.probs.data_ptr(),
                                          self.dlogits.data_ptr() if self.train else 0,
                                          self.frame_loss.data_ptr(), n, hw, 1.0 / ((global_batch or self.B) * self.T),
                                          LOSS_TYPES[loss_type], self._s())
        self.lib.rgp_reduce_sum(self.frame_loss.data_ptr(), n, self.loss.data_ptr(), self._s())
        return self.loss

    def read_state(self, t=None):
        t = self.T if t is None else t
        return self.Hf[t * self.B:(t + 1) * self.B, :self.H].contiguous()

    # ------------------------------------------------------------------ backward (BPTT) — base.py:278-281
    def _wgrad(self, Y, ldy, ycol0, Ny, X, rowsX, ldx, Nx, M, D):
        a = L.WgradArgs()
        a.Y, a.ldy, a.ycol0, a.Ny = Y, ldy, ycol0, Ny
        a.X, a.rowsX, a.ldx, a.xcol0, a.Nx = X, rowsX, ldx, 0, Nx
        a.M, a.ntaps = M, 1
        a.tap_off[0] = 0
        a.D, a.dtype, a.engine = D, self.dtype, self.engine
        self.lib.rgp_wgrad(C.byref(a), self._s())

    def backward(self):
        """Hand-written BPTT of the last forward(gt=...): fills self.grads (flat, op layouts).  dL/dlogits =
        (softmax - gt)/(B*T) was written by the fused softmax / loss kernel."""
        assert self.train, "construct the engine with train=True"
        lib, s = self.lib, self._s()
        B, T, P, Cp, Hp, Ip, H, Np, N = self.B, self.T, self.P, self.Cp, self.Hp, self.Ip, self.H, self.Np, self.N
        es = self.Hop.element_size()
        M = T * B
        g = lambda name: self._ptr(name, self.grads)
        lib.rgp_zero(self.grads.data_ptr(), self.grads.numel() * 4, s)
        # transposed operand copies of the weights (dX = dY . W for a forward Y = X . W^T)
        lib.rgp_transpose_taps(self._ptr("out_W_op"), self.outWT.data_ptr(), 1, Np, Hp, self.dtype, s)
        lib.rgp_transpose_taps(self._ptr("Whc_op"), self.WhcT.data_ptr(), 1, Hp, Hp, self.dtype, s)
        lib.rgp_transpose_taps(self._ptr("Whg_op"), self.WhgT.data_ptr(), 1, 2 * Hp, Hp, self.dtype, s)
        lib.rgp_transpose_taps(self._ptr("Wx_op"), self.WxT.data_ptr(), 1, 3 * Hp, Ip, self.dtype, s)
        # ---- output projection (gaze_rnn.py:346-348), all T at once
        lib.rgp_dense_dlogits(self.dlogits.data_ptr(), self.dL.data_ptr(), B, T, N, Np, self.dtype, s)
        h1 = self.Hop.data_ptr() + B * Hp * es                       # slab t+1 = h_{t+1}: the projection's input rows
        self._wgrad(self.dL.data_ptr(), Np, 0, Np, h1, self.Hop.shape[0] - B, Hp, Hp, M, g("out_W_op"))
        lib.rgp_colsum(self.dL.data_ptr(), Np, M, Np, g("out_b"), self.dtype, s)
        self._gemm(self.dL.data_ptr(), self.dL.shape[0], Np, self.outWT.data_ptr(), M, Hp, Np, self.dHhead.data_ptr(), Hp)
        # ---- the recurrence, backwards in time (tf.nn.rnn_cell.GRUCell)
        dh_rec, dh_next = 0, self.dhA
        for t in range(T - 1, -1, -1):
            sl = t * B * Hp * 4
            hprev = self.Hf.data_ptr() + sl                              # slab t = h_t, the step's input state
            dG_t = self.dG.data_ptr() + t * B * 3 * Hp * es
            lib.rgp_dense_gru_bwd_c(self.dHhead.data_ptr() + sl, dh_rec, self.Us.data_ptr() + sl, self.Cs.data_ptr() + sl,
                                    hprev, dG_t, dh_next.data_ptr(), B, H, Hp, self.dtype, s)
            self._gemm(dG_t + 2 * Hp * es, self.dG.shape[0] - t * B, 3 * Hp, self.WhcT.data_ptr(), B, Hp, Hp,
                       self.drh.data_ptr(), Hp)
            lib.rgp_dense_gru_bwd_r(self.drh.data_ptr(), self.Rs.data_ptr() + sl, hprev, dG_t, dh_next.data_ptr(), B, H, Hp,
                                    self.dtype, s)
            if t > 0:
                self._gemm(dG_t, self.dG.shape[0] - t * B, 3 * Hp, self.WhgT.data_ptr(), B, Hp, 2 * Hp,
                           dh_next.data_ptr(), Hp, accumulate=1)
            dh_rec = dh_next.data_ptr()
            dh_next = self.dhB if dh_next is self.dhA else self.dhA
        # ---- weight gradients of the cell: row contractions over all T*B rows
        self._wgrad(self.dG.data_ptr(), 3 * Hp, 0, 3 * Hp, self.X.data_ptr(), self.X.shape[0], Ip, Ip, M, g("Wx_op"))
        self._wgrad(self.dG.data_ptr(), 3 * Hp, 0, 2 * Hp, self.Hop.data_ptr(), self.Hop.shape[0], Hp, Hp, M, g("Whg_op"))
        self._wgrad(self.dG.data_ptr(), 3 * Hp, 2 * Hp, Hp, self.RH.data_ptr(), self.RH.shape[0], Hp, Hp, M, g("Whc_op"))
        lib.rgp_colsum(self.dG.data_ptr(), 3 * Hp, M, 2 * Hp, g("bias_g"), self.dtype, s)
        lib.rgp_colsum(self.dG.data_ptr() + 2 * Hp * es, 3 * Hp, M, Hp, g("bias_c"), self.dtype, s)
        # ---- input rows -> projection (gaze_rnn.py:294-304)
        self._gemm(self.dG.data_ptr(), self.dG.shape[0], 3 * Hp, self.WxT.data_ptr(), M, Ip, 3 * Hp, self.dX.data_ptr(), Ip)
        lib.rgp_dense_unpack(self.dX.data_ptr(), Ip, self.dPacc.data_ptr(), M, P, Cp, float(self.keep_prob),
                             int(self.dropout_seed), self.dtype, s)
        self._wgrad(self.dPacc.data_ptr(), Cp, 0, Cp, self.Apack.data_ptr(), self.Apack.shape[0], self.Cfeat, self.Cfeat,
                    M * P, g("proj_W_op"))
        lib.rgp_colsum(self.dPacc.data_ptr(), Cp, M * P, Cp, g("proj_b"), self.dtype, s)
