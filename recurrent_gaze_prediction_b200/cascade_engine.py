"""CascadeEngine — the two-level cascade model (BASELINE config 5; reference
``models/gaze_grcn_cascade.py:189-452``) over the C-ABI.  Forward / loss / inference.

Pipeline (one frame = one time step of one clip; frames are time-major, f = t*B + b):

  bottom level  (:257-312)  = :class:`GazeEngine` ("gaze_grcn", hidden 256, head "none"): projection 1024 -> 512, 3x3
                              ConvGRU over T steps; its bf16 / fp32 operand slabs h_1..h_T (8x8 p-layout rows per frame)
                              are consumed in place.
  up-sampling   (:314-336)  ONE 11x11 / stride-7 SAME transposed conv 7x7x256 -> 49x49x64 for all T frames = a 4-tap
                              GEMM over those rows (``RGP_EPI_UP7``): output pixel oy = 7i + ki - 2 gets h[i] with
                              ki = ry and h[i-1] with ki = ry + 7, i.e. block (a,b) of 7x7 pixels reads the 2x2 states
                              (a,b), (a-1,b), (a,b-1), (a-1,b-1) = row shifts {9, 8, 1, 0} in the p-layout; N = 49
                              pixels x 64 channels (+ padding to 13 x 256).  Written straight into the zero-bordered
                              operand layout of the top cell (51 x 51 rows per frame, data at (+2,+2)).
  top cell, input half (:346-351,379)  5x5 SAME conv 64 -> 3 gates x 3 units for all T frames = a 25-tap GEMM
                              (N = 16, K = 64; the resident-A-slab mode of the tcgen05 engine keeps the 128 + 208 rows
                              a tile touches in shared memory, so each input row is fetched once, not 25 times).
  top cell, recurrence       ``rgp_topgru_seq_fwd``: one persistent CTA per clip, state in shared memory.
  maxout FCs    (:383-420)  7203 -> 4802 -> max of halves 2401 -> 4802 -> max of halves 2401: two plain GEMMs over
                              all T*B rows + ``rgp_maxout``.  The halves tf.split separates are kept at columns
                              [0, 2401) and [2432, 4833) of a 4864-wide GEMM (19 x 256).
  prediction    (:422-424)  the 2401 values as a 49x49 map, NO softmax; loss (:427-441) = sum l2_loss / (B*T).

The shape bug of the file as written (top cell declared with 65 input channels, fed 64, :346-351,379) is resolved as in
``oracle/model_oracle.py``: reading A (default) the cell has 64 input channels; reading B (``top_in=65``) the
commented-out concat (:366-374) is live and the 65th channel is a frame-saliency map handed to ``forward(frm_sal=...)``
(the ShallowNet that would produce it is outside the path) - its 5x5 conv is evaluated inside the recurrence kernel.
Dropout on fc1 (:401-402) is live only while the reference trains; training of the cascade is not built.
"""
from __future__ import annotations

import ctypes as C
import os
from collections import OrderedDict

import numpy as np
import torch

from . import lib as L
from .engine import GazeEngine, LOSS_TYPES, _tdtype

GT = 49                      # top grid
TW = GT + 2                  # operand row width of the 5x5 cell (zero borders shared between rows)
TS = TW * TW                 # operand rows per frame (2601)
UPC = 64                     # up-sampled channels (:316)
UPN = 13 * 256               # 49 * 64 = 3136 GEMM columns padded to whole 256-wide tiles
NU = 3                       # units of the top cell (:346)
NFLAT = GT * GT * NU         # 7203
K1 = 7232                    # 7203 padded to a multiple of 64
NH = 2401                    # half of a 4802-wide maxout layer
NHP = 2432                   # ... padded to a multiple of 64 / 128
NFC = 2 * NHP                # 4864 = 19 * 256


def _pad(n, m):
    return (n + m - 1) // m * m


class CascadeEngine(object):
    model = "gaze_grcn_cascade"
    lstm = False
    use_bn = False
    G = 7
    GH = GT

    def __init__(self, B=2, T=35, Ch=256, Cin=512, Cfeat=1024, dtype=0, device="cuda", lib=None, train=False,
                 engine=None, top_in=64, seq=None, keep_top_states=False, fuse_top=None):
        if train:
            raise NotImplementedError("the cascade (gaze_grcn_cascade.py) is built for forward / loss / inference; its "
                                      "BPTT is not built")
        if top_in not in (UPC, UPC + 1):
            raise ValueError("top_in is 64 (reading A: the channels the cell is fed) or 65 (reading B: + frame saliency)")
        self.B, self.T, self.Ch, self.Cin, self.Cfeat = B, T, Ch, Cin, Cfeat
        self.dtype, self.train, self.top_in = dtype, False, top_in
        self.device = torch.device(device)
        self.lib = lib if lib is not None else L.get_lib()
        self.tdt = _tdtype(dtype)
        self.step = 0
        self.bottom = GazeEngine("gaze_grcn", B=B, T=T, G=7, Ch=Ch, Cin=Cin, Cfeat=Cfeat, dtype=dtype, device=device,
                                 lib=self.lib, head="none", engine=engine, seq=seq)
        self.engine = self.bottom.engine
        self.use_seq = self.bottom.use_seq
        self.keep_top_states = keep_top_states
        # up-sampling and the top cell's input convs as ONE composed GEMM (RGP_EPI_UP7X; tcgen05 engine only).  The
        # two-GEMM path stays for the fp32 engine and for reading the up-sampled maps (read_upsampled, net[...]).
        can_fuse = dtype == 1 and self.engine != 1 and hasattr(self.lib, "rgp_cascade_compose")
        if fuse_top is None:
            fuse_top = can_fuse and os.environ.get("RGP_CASCADE_FUSE", "1") != "0"
        if fuse_top and not can_fuse:
            raise ValueError("fuse_top needs the bf16 tcgen05 engine")
        self.fuse_top = bool(fuse_top)
        self._build_params()
        self._alloc()

    # ------------------------------------------------------------------ parameters of the top level
    def _build_params(self):
        spec = OrderedDict()
        spec["up_op"] = (4 * UPN, self.Ch)               # [tap (di,dj)][(ry*7+rx)*64 + co][ci]
        spec["Wtx_op"] = (25 * 16, UPC)                  # [tap ky*5+kx][gate*3 + unit (9 of 16)][ci]
        spec["uzr"] = (25, NU, 2 * NU)
        spec["uc"] = (25, NU, NU)
        spec["wsal"] = (25, 3 * NU)
        spec["fc1_op"] = (NFC, K1)
        spec["fc1_b"] = (NFC,)
        spec["fc2_op"] = (NFC, NHP)
        spec["fc2_b"] = (NFC,)
        self.spec, self.offsets = spec, OrderedDict()
        off = 0
        for k, shp in spec.items():
            n = int(np.prod(shp))
            self.offsets[k] = (off, n)
            off += _pad(n, 64)
        self.ntop = off
        self.tparams = torch.zeros(off, dtype=torch.float32, device=self.device)
        self.tparams_op = self.tparams if self.dtype == 0 else torch.zeros(off, dtype=self.tdt, device=self.device)

    @property
    def nparams(self):
        return self.bottom.nparams + self.ntop

    @property
    def params(self):
        """bottom-level and top-level flat parameter buffers, concatenated (a copy: checkpoints, inspection)"""
        return torch.cat([self.bottom.params, self.tparams])

    def _ptr(self, name, buf=None):
        off, _ = self.offsets[name]
        buf = self.tparams if buf is None else buf
        return buf.data_ptr() + off * buf.element_size()

    def _pop(self, name):
        return self._ptr(name, self.tparams_op)

    def view(self, name, buf=None):
        off, n = self.offsets[name]
        buf = self.tparams if buf is None else buf
        return buf[off:off + n].view(*self.spec[name])

    def load_tf_params(self, p):
        """Parameter dict with the reference's variable names (scope-qualified: RCNBottom/GRU_Conv_*, Upsampling/weight,
        RCNGaze/GRU_Conv_*, LastProjection/fc{1,2}/{weights,bias}) and TF layouts."""
        f = lambda k: np.asarray(p[k], dtype=np.float32)
        bot = {"proj_c3d_W": p["proj_c3d_W"], "proj_c3d_b": p["proj_c3d_b"]}
        for k, v in p.items():
            if k.startswith("RCNBottom/"):
                bot[k.split("/", 1)[1]] = v
        self.bottom.load_tf_params(bot)
        host = {k: np.zeros(s, dtype=np.float32) for k, s in self.spec.items()}
        # ---- up-sampling filter [11,11,64,Ch] (kH,kW,Cout,Cin) -> 4 tap matrices (gaze_grcn_cascade.py:317-331)
        wu = f("Upsampling/weight")
        assert wu.shape == (11, 11, UPC, self.Ch), wu.shape
        up = host["up_op"].reshape(4, UPN, self.Ch)
        for di in range(2):
            for dj in range(2):
                blk = np.zeros((7, 7, UPC, self.Ch), dtype=np.float32)
                ny, nx = (7 if di == 0 else 4), (7 if dj == 0 else 4)      # ki = ry + 7*di <= 10
                blk[:ny, :nx] = wu[7 * di:7 * di + ny, 7 * dj:7 * dj + nx]
                up[di * 2 + dj, :49 * UPC] = blk.reshape(49 * UPC, self.Ch)
        # ---- top cell (5x5, `top_in` input channels, 3 units)
        g = {nm: f("RCNGaze/GRU_Conv_" + nm) for nm in ("Wz", "Uz", "Wr", "Ur", "W", "U")}
        assert g["Wz"].shape == (5, 5, self.top_in, NU), (g["Wz"].shape, self.top_in)
        wx = np.concatenate([g["Wz"], g["Wr"], g["W"]], axis=3)                # [5,5,top_in,9]
        host["Wtx_op"].reshape(25, 16, UPC)[:, :3 * NU, :] = wx[:, :, :UPC, :].reshape(25, UPC, 3 * NU).transpose(0, 2, 1)
        if self.top_in == UPC + 1:
            host["wsal"][:] = wx[:, :, UPC, :].reshape(25, 3 * NU)
        host["uzr"][:] = np.concatenate([g["Uz"], g["Ur"]], axis=3).reshape(25, NU, 2 * NU)
        host["uc"][:] = g["U"].reshape(25, NU, NU)
        # ---- maxout FCs: [in, 4802] -> rows (outputs) of the two halves at 0 and NHP, K contiguous
        for nm, kin in (("fc1", NFLAT), ("fc2", NH)):
            W, b = f("LastProjection/%s/weights" % nm), f("LastProjection/%s/bias" % nm)
            assert W.shape == (kin, 2 * NH) and b.shape == (2 * NH,), (W.shape, b.shape)
            op = host[nm + "_op"]
            op[:NH, :kin] = W[:, :NH].T
            op[NHP:NHP + NH, :kin] = W[:, NH:].T
            host[nm + "_b"][:NH], host[nm + "_b"][NHP:NHP + NH] = b[:NH], b[NH:]
        flat = np.zeros(self.ntop, dtype=np.float32)
        for k, v in host.items():
            off, n = self.offsets[k]
            flat[off:off + n] = v.reshape(-1)
        self.tparams.copy_(torch.from_numpy(flat))
        self.refresh_operands()

    def op_to_tf(self, flat):
        """Inverse of load_tf_params on a flat host array laid out like `params` (bottom | top)."""
        flat = np.asarray(flat)
        nb = self.bottom.nparams
        out = OrderedDict()
        for k, v in self.bottom.op_to_tf(flat[:nb]).items():
            out[("RCNBottom/" + k) if k.startswith("GRU_Conv_") else k] = v
        top = flat[nb:]
        g = lambda k: top[self.offsets[k][0]:self.offsets[k][0] + self.offsets[k][1]].reshape(self.spec[k])
        up = g("up_op").reshape(4, UPN, self.Ch)
        wu = np.zeros((11, 11, UPC, self.Ch), dtype=flat.dtype)
        for di in range(2):
            for dj in range(2):
                ny, nx = (7 if di == 0 else 4), (7 if dj == 0 else 4)
                blk = up[di * 2 + dj, :49 * UPC].reshape(7, 7, UPC, self.Ch)
                wu[7 * di:7 * di + ny, 7 * dj:7 * dj + nx] = blk[:ny, :nx]
        out["Upsampling/weight"] = wu
        wx = np.zeros((5, 5, self.top_in, 3 * NU), dtype=flat.dtype)
        wx[:, :, :UPC, :] = g("Wtx_op").reshape(25, 16, UPC)[:, :3 * NU, :].transpose(0, 2, 1).reshape(5, 5, UPC, 3 * NU)
        if self.top_in == UPC + 1:
            wx[:, :, UPC, :] = g("wsal").reshape(5, 5, 3 * NU)
        uzr, uc = g("uzr").reshape(5, 5, NU, 2 * NU), g("uc").reshape(5, 5, NU, NU)
        for nm, w in (("Wz", wx[..., :NU]), ("Uz", uzr[..., :NU]), ("Wr", wx[..., NU:2 * NU]), ("Ur", uzr[..., NU:]),
                      ("W", wx[..., 2 * NU:]), ("U", uc)):
            out["RCNGaze/GRU_Conv_" + nm] = np.ascontiguousarray(w)
        for nm, kin in (("fc1", NFLAT), ("fc2", NH)):
            op, b = g(nm + "_op"), g(nm + "_b")
            out["LastProjection/%s/weights" % nm] = np.concatenate([op[:NH, :kin].T, op[NHP:NHP + NH, :kin].T], axis=1)
            out["LastProjection/%s/bias" % nm] = np.concatenate([b[:NH], b[NHP:NHP + NH]])
        return out

    def refresh_operands(self):
        if self.dtype != 0:
            self.lib.rgp_cast(self.tparams.data_ptr(), self.tparams_op.data_ptr(), self.ntop, self.dtype, self._s())
        if self.fuse_top:
            if getattr(self, "Wc", None) is None:
                self.Wc = torch.zeros(81 * 1024, self.Ch, dtype=self.tdt, device=self.device)
            self.lib.rgp_cascade_compose(self._ptr("up_op"), UPN, self._ptr("Wtx_op"), self.Wc.data_ptr(), self.Ch,
                                         self.dtype, self._s())

    # ------------------------------------------------------------------ buffers
    def _alloc(self):
        B, T, dev = self.B, self.T, self.device
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=dev)
        # up-sampled maps in the top cell's operand layout; +512 rows: the last tile's slab (128 + 4*51+4 rows) stays inside
        self.U = z(T * B * TS + 512, UPC, dt=self.tdt)
        self.Xg2 = z(T * B * TS, 16)
        rows = _pad(T * B, 128) + 128
        self.OG = z(rows, K1, dt=self.tdt)
        self.Htop = z(T * B, NFLAT) if self.keep_top_states else None
        self.top_state = z(B, NFLAT)
        self.Y = z(T * B, NFC)
        self.M1 = z(rows, NHP, dt=self.tdt)
        self.logits = z(B, T, GT, GT)          # the predicted maps (no softmax in this model: :422-424)
        self.probs = self.logits
        self.frame_loss = z(B * T)
        self.loss = z(1)

    def _s(self):
        if self.device.type == "cuda":
            return torch.cuda.current_stream(self.device).cuda_stream
        return 0

    def _gemm(self, **kw):
        a = L.TapGemmArgs()
        a.dtype, a.engine = self.dtype, self.engine
        taps = kw.pop("taps")
        a.ntaps = len(taps)
        for i, o in enumerate(taps):
            a.tap_off[i] = o
        for k, v in kw.items():
            setattr(a, k, v)
        self.lib.rgp_tapgemm(C.byref(a), self._s())

    def reset_state(self):
        self.bottom.reset_state()

    # ------------------------------------------------------------------ forward
    def forward(self, c3d=None, emb=None, gt=None, loss_type="l2", global_batch=None, carry_state=False, t0=0,
                frm_sal=None):
        """c3d [B,T,1024,7,7] -> predicted maps [B,T,49,49]; with gt also the l2 loss (gaze_grcn_cascade.py:427-441).
        frm_sal [B,T,49,49] fp32: the frame-saliency channel of reading B (required iff top_in == 65)."""
        assert emb is None and not carry_state and t0 == 0, "the cascade takes C3D features from h_0 = 0"
        assert (frm_sal is not None) == (self.top_in == UPC + 1), "frm_sal goes with top_in = 65"
        self.stage_bottom(c3d)
        if self.fuse_top:
            self.stage_top_fused()
        else:
            self.stage_upsample()
            self.stage_top_i2h()
        self.stage_top_recurrence(frm_sal)
        self.stage_fc()
        self.entry = "c3d"
        if gt is not None:
            self.loss_fwd(gt, loss_type, global_batch)
        return self.logits

    def stage_bottom(self, c3d):
        b = self.bottom
        if b._state_dirty:
            b.reset_state()
        b.t0 = 0
        b.stage_embed(c3d)
        b.stage_i2h()
        b.stage_recurrence()

    def stage_upsample(self):
        b, B, T = self.bottom, self.B, self.T
        taps = [9 - 8 * di - dj for di in range(2) for dj in range(2)]
        self._gemm(A=b._slab(b.Hop, 1), rowsA=b.Hop.shape[0] - B * b.S, lda=self.Ch, Bw=self._pop("up_op"),
                   M=T * B * b.S, N=UPN, K=self.Ch, taps=taps, epi=L.EPI_UP7, S=b.S, Wp=b.Wp, G=7, So=TS, Wo=TW, P=UPC,
                   Ch=self.Ch, B=B, T=T, out_op=self.U.data_ptr(), ldo=UPC)

    def stage_top_i2h(self):
        taps = [ky * TW + kx for ky in range(5) for kx in range(5)]
        self._gemm(A=self.U.data_ptr(), rowsA=self.U.shape[0], lda=UPC, Bw=self._pop("Wtx_op"), M=self.T * self.B * TS,
                   N=16, K=UPC, taps=taps, epi=L.EPI_STORE_F32, out_f32=self.Xg2.data_ptr(), ldo=16, S=TS, Wp=TW, G=GT)

    def stage_top_fused(self):
        """gaze_grcn_cascade.py:326-331 followed by the input convs of the top cell (:346-351, fed at :379) as one
        3x3-tap GEMM over the 7x7 grid with composed weights (include/rgp.h: RGP_EPI_UP7X, rgp_cascade_compose):
        4x fewer FLOPs than the two GEMMs and no 49x49x64 intermediate."""
        b, B, T = self.bottom, self.B, self.T
        self._gemm(A=b._slab(b.Hop, 1), rowsA=b.Hop.shape[0] - B * b.S, lda=self.Ch, Bw=self.Wc.data_ptr(),
                   M=T * B * 49, N=1024, K=self.Ch, taps=b.taps3, epi=L.EPI_UP7X, S=b.S, Wp=b.Wp, G=7, So=TS, Wo=TW,
                   out_f32=self.Xg2.data_ptr(), ldo=16, im2col=3)

    def stage_top_recurrence(self, frm_sal=None):
        sal = wsal = 0
        if frm_sal is not None:
            assert tuple(frm_sal.shape) == (self.B, self.T, GT, GT) and frm_sal.dtype == torch.float32
            sal, wsal = frm_sal.contiguous().data_ptr(), self._ptr("wsal")
        self.lib.rgp_topgru_seq_fwd(self.Xg2.data_ptr(), 16, TS, TW, self._ptr("uzr"), self._ptr("uc"), sal, wsal,
                                    self.top_state.data_ptr(), 0, self.OG.data_ptr(), K1,
                                    self.Htop.data_ptr() if self.Htop is not None else 0, self.B, self.T, self.dtype,
                                    self._s())

    def stage_fc(self):
        B, T, lib, s = self.B, self.T, self.lib, self._s()
        fc = dict(M=T * B, N=NFC, taps=[0], epi=L.EPI_STORE_F32, out_f32=self.Y.data_ptr(), ldo=NFC, S=1, Wp=1, G=1)
        self._gemm(A=self.OG.data_ptr(), rowsA=self.OG.shape[0], lda=K1, Bw=self._pop("fc1_op"), K=K1, **fc)
        lib.rgp_maxout(self.Y.data_ptr(), NFC, self._ptr("fc1_b"), NHP, NH, self.M1.data_ptr(), NHP, 0, B, T,
                       self.dtype, s)
        self._gemm(A=self.M1.data_ptr(), rowsA=self.M1.shape[0], lda=NHP, Bw=self._pop("fc2_op"), K=NHP, **fc)
        lib.rgp_maxout(self.Y.data_ptr(), NFC, self._ptr("fc2_b"), NHP, NH, 0, 0, self.logits.data_ptr(), B, T,
                       self.dtype, s)

    def loss_fwd(self, gt, loss_type="l2", global_batch=None):
        if loss_type != "l2":
            raise NotImplementedError("the cascade's loss is the summed l2 loss (gaze_grcn_cascade.py:427-441), not %r"
                                      % (loss_type,))
        assert tuple(gt.shape) == tuple(self.logits.shape) and gt.dtype == torch.float32
        n, hw = self.B * self.T, GT * GT
        self.lib.rgp_softmax_xent_fwd_bwd(self.logits.data_ptr(), gt.data_ptr(), 0, 0, self.frame_loss.data_ptr(), n, hw,
                                          1.0 / ((global_batch or self.B) * self.T), LOSS_TYPES["l2"], self._s())
        self.lib.rgp_reduce_sum(self.frame_loss.data_ptr(), n, self.loss.data_ptr(), self._s())
        return self.loss

    # ------------------------------------------------------------------ inspection (tests, net[...] of the model class)
    def read_state(self, t=None):
        """bottom state h_t [B,7,7,Ch] fp32"""
        return self.bottom.read_state(t)

    def read_upsampled(self, t):
        """up-sampled map of step t [B,49,49,64] (operand dtype -> fp32)"""
        if self.fuse_top:
            self.stage_upsample()        # the fused path never forms them
        n = self.B * TS
        u = self.U[t * n:(t + 1) * n].view(self.B, TW, TW, UPC)
        return u[:, 2:, 2:, :].float().contiguous()

    def read_top_state(self, t):
        """top cell output of step t [B,49,49,3] fp32 (needs keep_top_states=True)"""
        assert self.Htop is not None
        return self.Htop[t * self.B:(t + 1) * self.B].view(self.B, GT, GT, NU)

    def make_graph(self, fn):
        return self.bottom.make_graph(fn)

    def time_stages(self, c3d, gt, iters=3):
        names = ["bottom", "upsample", "top_i2h", "top_recurrence", "fc", "loss"]
        fns = [lambda: self.stage_bottom(c3d), self.stage_upsample, self.stage_top_i2h, self.stage_top_recurrence,
               self.stage_fc, lambda: self.loss_fwd(gt)]
        if self.fuse_top:
            names[1:3] = ["upsample+top_i2h (one composed GEMM)"]
            fns[1:3] = [self.stage_top_fused]
        acc = dict((n, 0.0) for n in names)
        for _ in range(iters):
            evs = [torch.cuda.Event(enable_timing=True) for _ in range(len(fns) + 1)]
            evs[0].record()
            for i, f in enumerate(fns):
                f()
                evs[i + 1].record()
            torch.cuda.synchronize()
            for i, n in enumerate(names):
                acc[n] += evs[i].elapsed_time(evs[i + 1]) / iters
        return acc
