"""GazeEngine — host orchestration of the RGP hot path over the C-ABI of ``librgp_sm100.so``.

One engine instance = one model replica on one GPU: it owns the flat parameter / gradient /
Adam buffers, the activation slabs, and issues the kernels of one forward (and backward) pass on
the current CUDA stream.  PyTorch is used for device allocations and streams only; all arithmetic
happens in the library (``include/rgp.h``).  Reference call sites are cited per step.

Row layouts: frames are time-major, f = t*B + b.  A G x G frame occupies
S = (G+1)^2 rows of C-channel vectors; position (y,x) is stored at "p-row" f*S+(y+1)*Wp+(x+1) of
an otherwise-zero operand buffer and its conv output lands at "q-row" f*S+y*Wp+x, which makes
every tap of a 3x3 SAME convolution a constant row shift ky*Wp+kx.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from collections import OrderedDict

import numpy as np
import torch

from . import lib as L

BN_EPS = 1e-3
HEAD_S, HEAD_W, HEAD_PAD = 100, 10, 3   # head-input layout: 10x10 rows per frame, data at (+3,+3)
LOSS_TYPES = {"xentropy": 0, "l2": 1, "KLD": 2}


def _tdtype(dtype):
    return torch.float32 if dtype == 0 else torch.bfloat16


class StepGraph(object):
    """One batch through the hot path as ONE CUDA-graph submission: forward + loss, and for train=True also BPTT +
    global-norm clip + TF-Adam with the schedule evaluated on the device (rgp_adam_schedule), so every replay is a correct
    optimiser step.  With a data-parallel `allreduce` the collective stays outside the captured region:
    graph(fwd + BPTT) -> NCCL all-reduce of the flat gradient -> graph(clip + Adam)   (base.py:287-297: the norm is of the
    GLOBAL gradient).  `fwd` is a callable issuing the forward on the engine's fixed input buffers.
    The warm-up run that precedes the capture is undone (parameters, moments and global_step restored), so building the
    graph does not train."""

    def __init__(self, eng, fwd, train=False, allreduce=None, opt=None):
        self.eng, self.train, self.allreduce = eng, bool(train), allreduce
        opt = dict(opt or {})
        opt.pop("allreduce", None)

        def fwd_bwd():
            fwd()
            if train:
                eng.backward()

        def update():
            eng.optimizer_step(**opt)

        def whole():
            fwd_bwd()
            update()

        snap = eng.snapshot_train_state() if train else None
        self.g2 = None
        if train and allreduce is not None:
            self.g1, n1 = eng.make_graph(fwd_bwd)
            self.g2, n2 = eng.make_graph(update)
            self.kernels = n1 + n2
        else:
            self.g1, self.kernels = eng.make_graph(whole if train else fwd_bwd)
        if snap is not None:
            eng.restore_train_state(snap)

    def replay(self):
        self.g1.replay()
        if self.g2 is not None:
            self.allreduce(self.eng.grads)
            self.g2.replay()
        if self.train:
            self.eng.note_replayed_steps(1)


class TrainStateMixin(object):
    """Flat-buffer training state shared by the engines (GazeEngine, DenseGruEngine): gradient / Adam-moment buffers with
    the parameter buffer's offsets, the device-side optimiser schedule, clip + TF-Adam (base.py:262-297,
    gaze_rnn.py:435-444), CUDA-graph capture of a step and the snapshot that undoes a capture's warm-up run.  Expects
    self.params / params_op / nparams / dtype / device / lib / train / param_version and self._s()."""

    def _alloc_train_state(self):
        dev, off = self.device, self.nparams
        self.grads = torch.zeros(off, dtype=torch.float32, device=dev)
        self.adam_m = torch.zeros(off, dtype=torch.float32, device=dev)
        self.adam_v = torch.zeros(off, dtype=torch.float32, device=dev)
        self.sumsq = torch.zeros(1, dtype=torch.float64, device=dev)
        # optimiser schedule state ON THE DEVICE (rgp_adam_schedule): global_step, initial learning rate, and the
        # (lr_t, lr) pair the Adam kernel reads — a captured graph of the training step replays a correct schedule
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.lr0_dev = torch.zeros(1, dtype=torch.float64, device=dev)
        self.sched = torch.zeros(2, dtype=torch.float32, device=dev)
        self._lr0_host = None
        self._step = 0
        self.optimizer = "adam"

    def set_optimizer(self, method):
        """models/base.py:268-275: 'adam' | 'rmsprop' (RMSProp, momentum 0.9) | 'sgd' (Momentum 0.9); anything else raises
        the reference's ValueError.  The two slot buffers are reused: adam_m = accumulator / momentum slot, adam_v = the
        RMSProp mean-square slot, which TF creates filled with ONES.  Call before the first step (it resets the slots)."""
        if method not in ("adam", "rmsprop", "sgd"):
            raise ValueError('Invalid optimization method!')
        self.optimizer = method
        self.adam_m.zero_()
        self.adam_v.fill_(1.0 if method == "rmsprop" else 0.0)

    def make_graph(self, fn):
        """Capture the launches fn() issues into a CUDA graph (one submission per step instead of ~80
        launches: the recurrence is launch-bound at small batch).  Returns (graph, kernels_per_replay)."""
        dev = self.device
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            fn()      # warm-up outside capture: function attributes, tensor-map cache, head composition
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        n0 = self.lib.launches
        with torch.cuda.graph(g):
            fn()
        return g, self.lib.launches - n0

    def snapshot_train_state(self):
        return (self.params.clone(), self.params_op.clone(), self.adam_m.clone(), self.adam_v.clone(), self._step,
                self.param_version)

    def restore_train_state(self, snap):
        p, po, m, v, step, ver = snap
        self.params.copy_(p)
        if self.params_op is not self.params:
            self.params_op.copy_(po)
        self.adam_m.copy_(m)
        self.adam_v.copy_(v)
        self.step = step
        self.param_version = ver + 1          # anything derived from the parameters (composed head) is stale

    # ------------------------------------------------------------------ optimiser — base.py:262-297, gaze_rnn.py:435-444
    @property
    def step(self):
        """global_step = optimiser updates applied so far.  The authoritative copy lives on the device (step_dev,
        advanced by rgp_adam_schedule inside the — possibly graph-captured — step); this host mirror is advanced by
        optimizer_step() / note_replayed_steps() and never needs a device read."""
        return getattr(self, "_step", 0)

    @step.setter
    def step(self, v):
        self._step = int(v)
        if self.train:
            self.step_dev.fill_(int(v))

    def step_on_device(self):
        """the device's own global_step (synchronises; tests / checkpoint sanity)"""
        return int(self.step_dev.cpu()[0])

    def note_replayed_steps(self, n=1):
        """a captured graph containing optimizer_step() was replayed n times: advance the host mirror"""
        self._step += int(n)
        self.param_version += int(n)

    def learning_rate(self, lr0=1e-4, decay=0.8, decay_steps=500):
        return lr0 * decay ** (self.step // decay_steps)

    def set_lr0(self, lr0):
        """initial learning rate (x decay_learning_rate's factor) -> the device scalar the schedule kernel reads; a
        host-to-device write only when the value changes (never inside a graph capture)"""
        if self._lr0_host != float(lr0):
            if self.device.type == "cuda" and torch.cuda.is_current_stream_capturing():
                raise RuntimeError("the learning rate changed inside a CUDA-graph capture: call set_lr0() before capturing")
            self.lr0_dev.fill_(float(lr0))
            self._lr0_host = float(lr0)

    def optimizer_step(self, lr0=1e-4, decay=0.8, clip=10.0, beta1=0.9, beta2=0.999, eps=1e-8, allreduce=None,
                       decay_steps=500):
        """clip_by_global_norm + TF-Adam on the flat buffers.  `allreduce(grads)` (optional) sums the gradient
        over data-parallel ranks first: the norm must be taken over the GLOBAL gradient (base.py:287-297).
        The staircase learning rate and Adam's bias-corrected step size are formed on the device from the device-side
        global_step (rgp_adam_schedule), so this call is graph-capturable with a correct schedule."""
        lib, s = self.lib, self._s()
        if allreduce is not None:
            allreduce(self.grads)
        self.set_lr0(lr0)
        lib.rgp_zero(self.sumsq.data_ptr(), 8, s)
        lib.rgp_sumsq(self.grads.data_ptr(), self.nparams, self.sumsq.data_ptr(), s)
        lib.rgp_adam_schedule(self.step_dev.data_ptr(), self.lr0_dev.data_ptr(), float(decay), int(decay_steps),
                              float(beta1), float(beta2), self.sched.data_ptr(), s)
        p_op = self.params_op.data_ptr() if self.dtype != 0 else 0
        if self.optimizer == "adam":
            lib.rgp_gradnorm_clip_adam_dev(self.params.data_ptr(), self.grads.data_ptr(), self.adam_m.data_ptr(),
                                           self.adam_v.data_ptr(), self.nparams, self.sumsq.data_ptr(), clip,
                                           self.sched.data_ptr(), beta1, beta2, eps, p_op, self.dtype, s)
        else:       # base.py:270-273: RMSProp(momentum 0.9; decay 0.9, epsilon 1e-10 = TF1 defaults) / Momentum(0.9)
            lib.rgp_gradnorm_clip_momentum_dev(self.params.data_ptr(), self.grads.data_ptr(), self.adam_m.data_ptr(),
                                               self.adam_v.data_ptr(), self.nparams, self.sumsq.data_ptr(), clip,
                                               self.sched.data_ptr(), 1 if self.optimizer == "rmsprop" else 0, 0.9, 0.9,
                                               1e-10, p_op, self.dtype, s)
        if not (self.device.type == "cuda" and torch.cuda.is_current_stream_capturing()):
            self._step += 1
            self.param_version += 1


class GazeEngine(TrainStateMixin):
    def __init__(self, model="gaze_grcn", B=2, T=35, G=7, Ch=128, Cin=512, Cfeat=1024, dtype=0,
                 device="cuda", lib=None, train=False, head=None, engine=None, seq=None, nsplit=None, mcast=None,
                 T_params=None, ntile=None):
        if model not in ("gaze_grcn", "gaze_grcn77", "gaze_lstm"):
            raise ValueError("unknown model %r" % (model,))
        self.model, self.B, self.T, self.G, self.Ch, self.Cin, self.Cfeat = model, B, T, G, Ch, Cin, Cfeat
        # Chunked long clips (extract_map.py:65,148-238 runs 105-step clips): the engine processes T frames per pass,
        # the model owns T_params >= T per-time-step parameter sets (the frozen-BN (gamma_t, beta_t) of
        # gaze_grcn.py:325); `t0` is the clip time of the chunk's first frame.  Inference only.
        self.Tp = T_params or T
        self.t0 = 0
        if self.Tp != T and train:
            raise ValueError("chunked time (T_params != T) is an inference feature; BPTT needs the whole clip")
        if engine is None:   # RGP_ENGINE=1 forces the FMA engine for bf16 operands (debug / A-B checks)
            engine = int(os.environ.get("RGP_ENGINE", "0"))
        self.dtype, self.train, self.engine = dtype, train, engine
        self.device = torch.device(device)
        self.lib = lib if lib is not None else L.get_lib()
        self.lstm = model == "gaze_lstm"
        # head: "deconv" (gaze_grcn.py:292-366), "fc" (gaze_grcn77.py:206-211) or "none" (the cascade's bottom level,
        # gaze_grcn_cascade.py:286-312: the states h_1..h_T are the product, CascadeEngine consumes them)
        self.head = head or ("fc" if model == "gaze_grcn77" else "deconv")
        if self.head not in ("deconv", "fc", "none"):
            raise ValueError("unknown head %r" % (self.head,))
        if self.head == "deconv" and G != 7:
            raise ValueError("the transposed-conv head (gaze_grcn.py:326-358) is defined for a 7x7 grid only")
        self.use_bn = (model == "gaze_grcn" and self.head == "deconv")
        self.GH = 49 if self.head == "deconv" else G
        self.Wp = G + 1
        self.S = self.Wp * self.Wp
        self.P = G * G
        self.tail = ((2 * self.Wp + 2 + 31) // 32) * 32
        self.ngx = 4 if self.lstm else 3        # gates fed by the input conv
        self.tdt = _tdtype(dtype)
        self.taps3 = [ky * self.Wp + kx for ky in range(3) for kx in range(3)]
        # Persistent ConvGRU recurrence (rgp_convgru_seq_fwd: all T steps in one launch, one CTA per 128-row tile).
        # Measured on B200: with the resident state slab it beats two tap-GEMM launches per step from 64 tiles up (128 clips,
        # 2-CTA channel split: 2.18 vs 2.81 ms; 256 clips: 2.63 vs 3.9 ms; profiles/r01_notes.md) and, at hidden 256 with
        # the channel split widened to 4 CTAs per tile, at every smaller batch too (profiles/r02_notes.md pass 49: 16 / 32 /
        # 64 / 96 clips: recurrence 2.02 / 2.03 / 2.08 / 2.38 ms per-step against 1.78 / 1.79 / 1.80 / 1.88 ms).  Hidden 128
        # likewise beyond the co-resident small-batch kernel's range (48 / 96 clips: 1.55 / 1.57 ms per-step against 0.98 / 0.99 ms
        # with 64-channel slices).  RGP_SEQ=0/1 forces the choice
        # (A/B checks); nsplit / mcast are the kernel's cluster options.
        mtiles = (B * self.S + 127) // 128
        seq_arg = seq
        gfits = getattr(self.lib, "rgp_convgru_seq_small_fits", None)
        if seq is None:
            env = os.environ.get("RGP_SEQ", "")
            # the co-resident small-batch kernel (below) is the faster one where it fits (hidden <= 128, <= 36 clips)
            small_ok = bool((not self.lstm) and engine != 1 and dtype == 1 and gfits is not None and gfits(B, G, Ch) and
                            env != "0")
            seq = (dtype == 1 and engine != 1 and Ch in (128, 256) and (env == "1" or (env != "0" and not small_ok)))
        self.use_seq = bool(seq) and seq != "small" and (not self.lstm) and 128 % self.S == 0
        # Small-batch persistent ConvGRU (rgp_convgru_seq_small_fwd: weight slices resident in shared memory, two counter
        # exchanges per step) - the regime below gru_seq's 64 tiles where all tiles x Ch/16 CTAs are co-resident (hidden 128:
        # up to 36 clips at 7x7, the reference's batch 28).
        self.use_seq_small = bool((not self.lstm) and (not self.use_seq) and seq_arg is not False and engine != 1 and
                                  gfits is not None and (dtype == 1 or seq_arg == "small") and gfits(B, G, Ch) and
                                  os.environ.get("RGP_SEQ", "") != "0")
        # Persistent ConvLSTM recurrence (rgp_convlstm_seq_fwd: all T steps in one launch, any grid; W_h slices resident in
        # shared memory, tiles exchange their halo rows through per-tile counters).  Taken whenever all its CTAs
        # (tiles x Ch/16) are co-resident - the small-batch, latency-bound regime (14x14: up to 10 clips; 7x7: up to 36).
        fits = getattr(self.lib, "rgp_convlstm_seq_fits", None)
        self.use_seq_lstm = bool(self.lstm and seq_arg is not False and engine != 1 and fits is not None and
                                 (dtype == 1 or seq_arg is True) and fits(B, G, Ch) and
                                 os.environ.get("RGP_SEQ", "") != "0")
        # channel split at hidden 256: as many CTAs per tile (4, 2, 1) as keeps tiles x nsplit within one wave of 148 SMs
        # (64-channel slices at the finest: 4-way at hidden 256, 2-way at hidden 128)
        self.nsplit = nsplit or int(os.environ.get("RGP_SEQ_NSPLIT", "0")) or \
            ((4 if mtiles <= 37 else 2 if mtiles <= 74 else 1) if Ch == 256 else (2 if (Ch == 128 and mtiles <= 74) else 1))
        # Two tiles per CTA (ntile = 2, with the 2-CTA channel split): every streamed weight box feeds the MMAs of two tiles,
        # halving the bytes an SM pulls through its L2->SMEM port per MMA.  Parity-tested; MEASURED SLOWER at the bench batch
        # (recurrence 2.98 vs 2.29 ms at 256 clips, gpurun_out/r02p11_*): the phases are not bound by the weight stream
        # alone, and the channel split adds two cross-CTA hand-offs per step.  Off by default (RGP_SEQ_NTILE=2 / ntile=2).
        self.ntile = ntile or int(os.environ.get("RGP_SEQ_NTILE", "0")) or 1
        if self.ntile == 2:
            self.nsplit = 2
        # mcast > 1: every weight box is fetched once per cluster and TMA-multicast; measured slower (the bound is the
        # SM's own L2->SMEM port, which multicast does not relieve), so the default is 1
        self.mcast = mcast or int(os.environ.get("RGP_SEQ_MCAST", "0")) or 1
        self._build_params()
        self._alloc()
        self.head_version = -1
        self.param_version = 0
        self._prof = None             # list of (section, cuda event) while time_backward() runs
        self._state_dirty = False     # slab 0 (h_0, and c_0 for the LSTM) holds a caller-provided state

    # ------------------------------------------------------------------ parameters
    def _build_params(self):
        Ch, Cin, T, G = self.Ch, self.Cin, self.T, self.G
        spec = OrderedDict()
        spec["proj_W_op"] = (Cin, self.Cfeat)
        spec["proj_b"] = (Cin,)
        spec["Wx_op"] = (9 * self.ngx * Ch, Cin)
        if self.lstm:
            spec["Wh_op"] = (9 * 3 * Ch, Ch)
            spec["Whc_dead"] = (3, 3, Ch, Ch)
            for nm in ("wci", "wcf", "wco"):
                spec[nm] = (G, G, Ch)
        else:
            spec["Uzr_op"] = (9 * 2 * Ch, Ch)
            spec["Uc_op"] = (9 * Ch, Ch)
        if self.head == "deconv":
            spec["w1"] = (5, 5, 64, Ch)
            spec["w2"] = (5, 5, 32, 64)
            spec["w3"] = (7, 7, 12, 32)
            spec["out_W"] = (12,)
        elif self.head == "fc":
            spec["out_W"] = (Ch,)
        if self.head != "none":
            spec["out_b"] = (1,)
        if self.use_bn:
            spec["bn_gamma"] = (self.Tp, Ch)
            spec["bn_beta"] = (self.Tp, Ch)
        self.spec = spec
        self.offsets = OrderedDict()
        off = 0
        for k, shp in spec.items():
            n = int(np.prod(shp))
            self.offsets[k] = (off, n)
            off += (n + 63) // 64 * 64      # keep every tensor 256-byte aligned (TMA / vector loads)
        self.nparams = off
        dev = self.device
        self.params = torch.zeros(off, dtype=torch.float32, device=dev)
        self.params_op = self.params if self.dtype == 0 else torch.zeros(off, dtype=self.tdt, device=dev)
        if self.train:
            self._alloc_train_state()

    def view(self, name, buf=None):
        off, n = self.offsets[name]
        buf = self.params if buf is None else buf
        return buf[off:off + n].view(*self.spec[name])

    def _ptr(self, name, buf=None):
        off, _ = self.offsets[name]
        buf = self.params if buf is None else buf
        return buf.data_ptr() + off * buf.element_size()

    def load_tf_params(self, p):
        """Load a parameter dict with the reference's variable names and TF layouts
        (conv kernels [kH,kW,Cin,Cout], conv_transpose kernels [kH,kW,Cout,Cin])."""
        Ch = self.Ch

        def conv_op(ws):  # list of [3,3,Cin,Cout] -> [9*sum(Cout), Cin]
            w = np.concatenate([np.asarray(x, dtype=np.float32) for x in ws], axis=3)  # [3,3,Cin,N]
            return np.ascontiguousarray(w.transpose(0, 1, 3, 2)).reshape(-1, w.shape[2])

        host = {}
        if "proj_c3d_W" in p:
            host["proj_W_op"] = np.ascontiguousarray(np.asarray(p["proj_c3d_W"], dtype=np.float32).T)
            host["proj_b"] = np.asarray(p["proj_c3d_b"], dtype=np.float32)
        if self.lstm:
            host["Wx_op"] = conv_op([p["ConvLSTM_Wx" + g] for g in "ifco"])
            host["Wh_op"] = conv_op([p["ConvLSTM_Wh" + g] for g in "ifo"])
            host["Whc_dead"] = np.asarray(p["ConvLSTM_Whc"], dtype=np.float32)
            for g in "ifo":
                host["wc" + g] = np.asarray(p["ConvLSTM_Wc" + g], dtype=np.float32)
        else:
            host["Wx_op"] = conv_op([p["GRU_Conv_Wz"], p["GRU_Conv_Wr"], p["GRU_Conv_W"]])
            host["Uzr_op"] = conv_op([p["GRU_Conv_Uz"], p["GRU_Conv_Ur"]])
            host["Uc_op"] = conv_op([p["GRU_Conv_U"]])
        if self.head == "deconv":
            host["w1"] = np.asarray(p["Upsampling/weight1"], dtype=np.float32)
            host["w2"] = np.asarray(p["Upsampling/weight2"], dtype=np.float32)
            host["w3"] = np.asarray(p["Upsampling/weight3"], dtype=np.float32)
        if self.head != "none":
            host["out_W"] = np.asarray(p["out_W"], dtype=np.float32).reshape(-1)
            host["out_b"] = np.asarray(p["out_b"], dtype=np.float32).reshape(-1)
        if self.use_bn:
            host["bn_gamma"] = np.asarray(p["bn_gamma"], dtype=np.float32)
            host["bn_beta"] = np.asarray(p["bn_beta"], dtype=np.float32)
        flat = np.zeros(self.nparams, dtype=np.float32)
        for k, v in host.items():
            off, n = self.offsets[k]
            assert v.size == n, (k, v.shape, self.spec[k])
            flat[off:off + n] = v.reshape(-1)
        self.params.copy_(torch.from_numpy(flat))
        self.refresh_operands()

    def op_to_tf(self, flat):
        """Inverse of load_tf_params on a flat host array (parameters or gradients)."""
        Ch, Cin = self.Ch, self.Cin
        flat = np.asarray(flat)

        def get(k):
            off, n = self.offsets[k]
            return flat[off:off + n].reshape(self.spec[k])

        def conv_tf(k, cin, names):
            w = get(k).reshape(3, 3, -1, cin).transpose(0, 1, 3, 2)   # [3,3,cin,N]
            return {nm: np.ascontiguousarray(w[..., i * Ch:(i + 1) * Ch]) for i, nm in enumerate(names)}

        out = OrderedDict()
        out["proj_c3d_W"] = np.ascontiguousarray(get("proj_W_op").T)
        out["proj_c3d_b"] = get("proj_b").copy()
        if self.lstm:
            out.update(conv_tf("Wx_op", Cin, ["ConvLSTM_Wx" + g for g in "ifco"]))
            out.update(conv_tf("Wh_op", Ch, ["ConvLSTM_Wh" + g for g in "ifo"]))
            out["ConvLSTM_Whc"] = get("Whc_dead").copy()
            for g in "ifo":
                out["ConvLSTM_Wc" + g] = get("wc" + g).copy()
        else:
            out.update(conv_tf("Wx_op", Cin, ["GRU_Conv_Wz", "GRU_Conv_Wr", "GRU_Conv_W"]))
            out.update(conv_tf("Uzr_op", Ch, ["GRU_Conv_Uz", "GRU_Conv_Ur"]))
            out.update(conv_tf("Uc_op", Ch, ["GRU_Conv_U"]))
        if self.head == "deconv":
            out["Upsampling/weight1"] = get("w1").copy()
            out["Upsampling/weight2"] = get("w2").copy()
            out["Upsampling/weight3"] = get("w3").copy()
            out["out_W"] = get("out_W").reshape(12, 1).copy()
        elif self.head == "fc":
            out["out_W"] = get("out_W").reshape(Ch, 1).copy()
        if self.head != "none":
            out["out_b"] = get("out_b").copy()
        if self.use_bn:
            out["bn_gamma"] = get("bn_gamma").copy()
            out["bn_beta"] = get("bn_beta").copy()
        return out

    def refresh_operands(self):
        """Re-derive everything that depends on the parameter values (operand-dtype copy, composed head)."""
        if self.dtype != 0:
            self.lib.rgp_cast(self.params.data_ptr(), self.params_op.data_ptr(), self.nparams, self.dtype, self._s())
        self.param_version += 1

    # ------------------------------------------------------------------ buffers
    def _alloc(self):
        B, T, S, Ch, Cin = self.B, self.T, self.S, self.Ch, self.Cin
        dev, tdt = self.device, self.tdt
        z = lambda *shape, dt=torch.float32: torch.zeros(*shape, dtype=dt, device=dev)
        self.rows_step = B * S + self.tail        # rows a step-GEMM may touch: its slab + the zero halo behind it
        self.Apack = z(T * B * self.P, self.Cfeat, dt=tdt)
        self.Xp = z(T * B * S + self.tail, Cin, dt=tdt)
        self.Xg = z(T * B * S, self.ngx * Ch)
        # h_0..h_T as contiguous B*S-row slabs (slab t = h_{t-1}): operand copy (p-layout) + fp32 master
        self.Hop = z((T + 1) * B * S + self.tail, Ch, dt=tdt)
        self.Hf = z((T + 1) * B * S + self.tail, Ch)
        nsl = T if self.train else 1
        if self.lstm:
            self.Cf = z(self.rows_step, Ch)
            self.Hconv = z(B * S, 3 * Ch)
            wsb = getattr(self.lib, "rgp_convlstm_seq_workspace_bytes", None)
            nflag = (wsb(B, self.G) // 4) if wsb is not None else (B * S + 127) // 128 + 1
            self.seq_flags = torch.zeros(nflag, dtype=torch.int32, device=dev)
            self.Lsave = z(T, B * S, 5 * Ch) if self.train else None
        else:
            self.RH = z(nsl * B * S + self.tail, Ch, dt=tdt)
            wsb = getattr(self.lib, "rgp_convgru_seq_small_workspace_bytes", None)
            nflag = (wsb(B, self.G) // 4) if wsb is not None else 2 * ((B * S + 127) // 128) + 1
            self.seq_flags = torch.zeros(nflag, dtype=torch.int32, device=dev)
            self.Z = z(nsl, B * S, Ch)
            self.Rs = z(T, B * S, Ch) if self.train else None
            self.Cs = z(T, B * S, Ch) if self.train else None
        hw = self.GH * self.GH
        if self.head == "deconv":
            self.HS = z(T * B * HEAD_S + 128, Ch, dt=tdt)
            self.w3f = z(49 * 32)
            self.w23 = z(121 * 64)
            self.w123 = z(361 * Ch)
            self.bhead = z(16 * 48, Ch, dt=tdt)
        self.logits = z(B, T, self.GH, self.GH)
        self.probs = z(B, T, self.GH, self.GH)
        self.frame_loss = z(B * T)
        self.loss = z(1)
        if self.train:
            self.dlogits = z(B, T, self.GH, self.GH)
            # dL/d(pre-activations) of the input-fed gates, columns z | r | c (GRU) or i | f | c | o (LSTM)
            self.dG = z(T * B * S, self.ngx * Ch, dt=tdt)
            self.dhs = z(T * B * S, Ch)                     # dL/d(head input) in the recurrence q-layout
            self.dhpart = z(B * S, Ch)                      # GRU: non-conv part of dL/dh_{t-1}; LSTM: dL/dc
            if self.lstm:
                self.dHc = z(T * B * S, 3 * Ch, dt=tdt)     # dL/d(hidden conv outputs), columns (i+c) | f | o
                self.dhrec = z(B * S, Ch)                   # conv^T term of dL/dh_{t-1}
                self.WhT = z(9 * Ch, 3 * Ch, dt=tdt)
            else:
                self.UcT = z(9 * Ch, Ch, dt=tdt)
                self.UzrT = z(9 * Ch, 2 * Ch, dt=tdt)
            self.WxT = z(9 * Cin, self.ngx * Ch, dt=tdt)
            self.dEmb = z(T * B * self.P, Cin, dt=tdt)
            if self.head == "deconv":
                self.patches = z(T * B * HEAD_S + 128, 384, dt=tdt)
                self.w123T = z(Ch, 384, dt=tdt)
                self.dw123 = z(384, Ch)
                self.dw23 = z(121 * 64)
                self.dw3f = z(49 * 32)

    def _slab(self, buf, t):
        """device address of the B*S-row slab t of a flat [rows, C] buffer"""
        return buf.data_ptr() + t * self.B * self.S * buf.shape[-1] * buf.element_size()

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

    # ------------------------------------------------------------------ forward
    def forward(self, c3d=None, emb=None, gt=None, loss_type="xentropy", global_batch=None, carry_state=False, t0=0):
        """c3d [B,T,Cfeat,G,G] fp32 or bf16 (reference entry, gaze_grcn.py:173) or emb [B,T,G,G,Cin] fp32
        (already-projected features).  Returns logits [B,T,GH,GW]; with gt also fills probs and loss.
        carry_state=True starts from the state left in slab 0 by set_initial_state() / carry_last_state()
        (cell-level API, chunked long clips) instead of h_0 = 0 (gaze_grcn.py:262)."""
        if not carry_state and self._state_dirty:
            self.reset_state()
        self.t0 = t0            # clip time of this chunk's first frame (selects the per-time-step BN parameters)
        self.stage_embed(c3d, emb)
        self.stage_i2h()
        self.stage_recurrence(zero_cell=not carry_state)
        self.stage_head()
        self.entry = "c3d" if c3d is not None else "emb"
        if gt is not None:
            self.loss_fwd(gt, loss_type, global_batch)
        return self.logits

    def _pop(self, name):
        return self._ptr(name, self.params_op)

    def stage_embed(self, c3d=None, emb=None):
        """(a1) layout + projection: gaze_grcn.py:223-254."""
        lib, s = self.lib, self._s()
        B, T, S, Cin, Wp, G = self.B, self.T, self.S, self.Cin, self.Wp, self.G
        if c3d is not None:
            # fp32 as the reference feeds them (gaze_rnn.py:520-531), or bf16: half the host->device bytes, and since the
            # bf16 engine rounds the features to bf16 in this very pass, the packed operand is bit-identical
            assert tuple(c3d.shape) == (B, T, self.Cfeat, G, G) and c3d.dtype in (torch.float32, torch.bfloat16)
            lib.rgp_pack_c3d_src(c3d.data_ptr(), 1 if c3d.dtype == torch.bfloat16 else 0, self.Apack.data_ptr(), B, T,
                                 self.Cfeat, self.P, self.dtype, s)
            self._gemm(A=self.Apack.data_ptr(), rowsA=T * B * self.P, lda=self.Cfeat, Bw=self._pop("proj_W_op"),
                       M=T * B * self.P, N=Cin, K=self.Cfeat, taps=[0], epi=L.EPI_PROJ, P=self.P, G=G,
                       So=S, Wo=Wp, pado=1, out_op=self.Xp.data_ptr(), ldo=Cin, bias=self._ptr("proj_b"))
        else:
            assert tuple(emb.shape) == (B, T, G, G, Cin) and emb.dtype == torch.float32
            lib.rgp_pack_emb(emb.data_ptr(), self.Xp.data_ptr(), B, T, G, Cin, self.dtype, s)

    def stage_i2h(self):
        """(a3) input half of every gate for all T frames at once: gaze_grcn.py:108-109,112-113,122-123."""
        NX = self.ngx * self.Ch
        # Dense-M mode on the tcgen05 engine: only the G*G real positions of a frame are GEMM rows (49 instead of the 64
        # p-layout rows at G = 7: -23 % MMA work), each tap's A tile comes from a TMA im2col load with hardware zero fill at
        # the frame borders; rows land at the same q-rows of Xg.  RGP_IM2COL=0 restores the slab path (A/B runs).
        dense = (self.dtype == 1 and self.engine != 1 and NX % 32 == 0 and os.environ.get("RGP_IM2COL", "1") != "0")
        self._gemm(A=self.Xp.data_ptr(), rowsA=self.Xp.shape[0], lda=self.Cin, Bw=self._pop("Wx_op"),
                   M=self.T * self.B * (self.P if dense else self.S), N=NX, K=self.Cin, taps=self.taps3,
                   epi=L.EPI_STORE_F32, out_f32=self.Xg.data_ptr(), ldo=NX, S=self.S, Wp=self.Wp, G=self.G,
                   im2col=1 if dense else 0)

    # ---- recurrent state access (zero_state / cell-level API: gaze_grcn.py:132-146, gaze_lstm.py:136-148)
    def _pview(self, buf, t):
        """slab t of a p-layout buffer as [B, Wp, Wp, C]; position (y, x) lives at [:, y+1, x+1]"""
        n = self.B * self.S
        return buf[t * n:(t + 1) * n].view(self.B, self.Wp, self.Wp, buf.shape[-1])

    def reset_state(self):
        n = self.B * self.S
        self.lib.rgp_zero(self.Hf.data_ptr(), n * self.Ch * 4, self._s())
        self.lib.rgp_zero(self.Hop.data_ptr(), n * self.Ch * self.Hop.element_size(), self._s())
        self._state_dirty = False

    def set_initial_state(self, h, c=None):
        """h (and c for the LSTM) [B,G,G,Ch] fp32 device tensors -> slab 0 (fp32 master + operand copy)."""
        B, G, Ch, s = self.B, self.G, self.Ch, self._s()
        assert tuple(h.shape) == (B, G, G, Ch) and h.dtype == torch.float32, "state must be [B,%d,%d,%d] fp32" % (G, G, Ch)
        h = h.contiguous()
        self.lib.rgp_pack_emb(h.data_ptr(), self.Hf.data_ptr(), B, 1, G, Ch, 0, s)
        self.lib.rgp_pack_emb(h.data_ptr(), self.Hop.data_ptr(), B, 1, G, Ch, self.dtype, s)
        if self.lstm:
            assert c is not None and tuple(c.shape) == (B, G, G, Ch)
            self.lib.rgp_zero(self.Cf.data_ptr(), self.Cf.numel() * 4, s)
            self.lib.rgp_pack_emb(c.contiguous().data_ptr(), self.Cf.data_ptr(), B, 1, G, Ch, 0, s)
        self._state_dirty = True

    def read_state(self, t=None):
        """h_t (after t steps; default: after the last step) as [B,G,G,Ch] fp32; LSTM: (c, h) with c the current cell."""
        t = self.T if t is None else t
        h = self._pview(self.Hf, t)[:, 1:, 1:, :].contiguous()
        if self.lstm:
            return self._pview(self.Cf, 0)[:, 1:, 1:, :].contiguous(), h
        return h

    def carry_last_state(self):
        """h_T -> slab 0, so that the next forward(carry_state=True) continues the same clips (chunked long clips,
        extract_map.py:148-238)."""
        n = self.B * self.S
        self.Hf[:n].copy_(self.Hf[self.T * n:(self.T + 1) * n])
        self.Hop[:n].copy_(self.Hop[self.T * n:(self.T + 1) * n])
        self._state_dirty = True

    def stage_recurrence(self, zero_cell=True):
        """(a4/a6/a7) the T-step recurrence: gaze_grcn.py:259-288, gaze_lstm.py:260-286."""
        lib, s = self.lib, self._s()
        B, T, S, Ch, Wp, G = self.B, self.T, self.S, self.Ch, self.Wp, self.G
        assert 0 <= self.t0 and self.t0 + T <= self.Tp, "chunk [t0, t0+T) outside the model's %d time steps" % self.Tp
        es = self.Hop.element_size()
        pop = self._pop
        NX = self.ngx * Ch
        M = B * S
        if self.use_seq or self.use_seq_small:
            a = L.GruSeqArgs()
            a.Hop, a.rowsH, a.Hf = self.Hop.data_ptr(), self.Hop.shape[0], self.Hf.data_ptr()
            a.RH, a.rowsRH = self.RH.data_ptr(), self.RH.shape[0]
            a.Uzr, a.Uc, a.Xg = pop("Uzr_op"), pop("Uc_op"), self.Xg.data_ptr()
            a.Zs = self.Z.data_ptr()
            if self.train:
                a.Rs, a.Cs = self.Rs.data_ptr(), self.Cs.data_ptr()
            if self.head == "deconv":
                a.HS = self.HS.data_ptr()
            if self.use_bn:
                a.bn_gamma, a.bn_beta = self._ptr("bn_gamma") + self.t0 * Ch * 4, self._ptr("bn_beta") + self.t0 * Ch * 4
                a.alpha = 1.0 / math.sqrt(1.0 + BN_EPS)
            a.B, a.T, a.G, a.Ch = B, T, G, Ch
            a.So, a.Wo, a.pado = HEAD_S, HEAD_W, HEAD_PAD
            a.nsplit, a.dtype, a.mcast, a.ntile = self.nsplit, self.dtype, self.mcast, self.ntile
            if self.use_seq_small:
                lib.rgp_zero(self.seq_flags.data_ptr(), self.seq_flags.numel() * 4, s)
                lib.rgp_convgru_seq_small_fwd(C.byref(a), self.seq_flags.data_ptr(), s)
            else:
                lib.rgp_convgru_seq_fwd(C.byref(a), s)
            if self.head == "fc":   # gaze_grcn77.py:206-211
                self._fc_head_all(s)
            return
        if self.use_seq_lstm:
            if zero_cell:
                lib.rgp_zero(self.Cf.data_ptr(), self.Cf.numel() * 4, s)
            lib.rgp_zero(self.seq_flags.data_ptr(), self.seq_flags.numel() * 4, s)
            a = L.LstmSeqArgs()
            a.Hop, a.rowsH, a.Hf, a.Cf = self.Hop.data_ptr(), self.Hop.shape[0], self.Hf.data_ptr(), self.Cf.data_ptr()
            a.Wh, a.Xg = pop("Wh_op"), self.Xg.data_ptr()
            a.wci, a.wcf, a.wco = self._ptr("wci"), self._ptr("wcf"), self._ptr("wco")
            a.HS = self.HS.data_ptr() if self.head == "deconv" else 0
            a.save = self.Lsave.data_ptr() if self.train else 0
            a.flags = self.seq_flags.data_ptr()
            a.B, a.T, a.G, a.Ch = B, T, G, Ch
            a.So, a.Wo, a.pado, a.dtype = HEAD_S, HEAD_W, HEAD_PAD, self.dtype
            lib.rgp_convlstm_seq_fwd(C.byref(a), s)
            if self.head == "fc":
                self._fc_head_all(s)
            return
        for t in range(T):
            xg = self.Xg.data_ptr() + t * M * NX * 4
            hs = 0
            if self.head == "deconv":
                hs = self.HS.data_ptr() + t * B * HEAD_S * Ch * es
            if self.lstm:
                if t == 0 and zero_cell:
                    lib.rgp_zero(self.Cf.data_ptr(), self.Cf.numel() * 4, s)
                self._gemm(A=self._slab(self.Hop, t), rowsA=self.rows_step, lda=Ch, Bw=pop("Wh_op"), M=M, N=3 * Ch,
                           K=Ch, taps=self.taps3, epi=L.EPI_STORE_F32, out_f32=self.Hconv.data_ptr(), ldo=3 * Ch,
                           S=S, Wp=Wp, G=G)
                lib.rgp_lstm_gates(xg, NX, self.Hconv.data_ptr(), self._ptr("wci"), self._ptr("wcf"),
                                   self._ptr("wco"), self.Cf.data_ptr(), self._slab(self.Hf, t + 1),
                                   self._slab(self.Hop, t + 1), hs, B, G, Ch, HEAD_S, HEAD_W, HEAD_PAD, self.dtype,
                                   self.Lsave[t].data_ptr() if self.train else 0, s)
            else:
                sl = t if self.train else 0
                common = dict(M=M, K=Ch, taps=self.taps3, S=S, Wp=Wp, G=G, Ch=Ch, B=B, T=T, t=t, xg=xg, ldx=NX,
                              rowsA=self.rows_step, lda=Ch, hprev_f32=self._slab(self.Hf, t),
                              z_f32=self.Z[sl].data_ptr())
                self._gemm(A=self._slab(self.Hop, t), Bw=pop("Uzr_op"), N=2 * Ch, epi=L.EPI_GRU_ZR,
                           out_op=self._slab(self.RH, sl), save_r=self.Rs[t].data_ptr() if self.train else 0,
                           **common)
                extra = {}
                if self.use_bn:
                    tb = self.t0 + t
                    extra = dict(bias=self._ptr("bn_gamma") + tb * Ch * 4, bias2=self._ptr("bn_beta") + tb * Ch * 4,
                                 alpha=1.0 / math.sqrt(1.0 + BN_EPS))
                self._gemm(A=self._slab(self.RH, sl), Bw=pop("Uc_op"), N=Ch, epi=L.EPI_GRU_C,
                           hnext_f32=self._slab(self.Hf, t + 1), out_op=self._slab(self.Hop, t + 1), out_op2=hs,
                           So=HEAD_S, Wo=HEAD_W, pado=HEAD_PAD, save_c=self.Cs[t].data_ptr() if self.train else 0,
                           **common, **extra)
        if self.head == "fc":   # gaze_grcn77.py:206-211
            self._fc_head_all(s)

    def seq_timed_out(self):
        """True if the last persistent ConvLSTM launch gave up waiting for a neighbouring tile (synchronises; the kernel
        never hangs: it reports through the last counter and the pass's results are then invalid)"""
        if self.use_seq_lstm:
            return int(self.seq_flags[-1].cpu()) != 0
        if getattr(self, "use_seq_small", False):
            return int(self.seq_flags[(self.B * self.S + 127) // 128].cpu()) != 0
        return False

    def _fc_head_all(self, s):
        """per-position FC head on the fp32 states of all T steps (they stay resident in Hf): one launch"""
        lib, B, T, G, Ch = self.lib, self.B, self.T, self.G, self.Ch
        if Ch % 4 == 0:
            lib.rgp_fc_head_all(self._slab(self.Hf, 1), self._ptr("out_W"), self._ptr("out_b"),
                                self.logits.data_ptr(), B, T, G, Ch, s)
        else:
            for t in range(T):
                lib.rgp_fc_head(self._slab(self.Hf, t + 1), self._ptr("out_W"), self._ptr("out_b"),
                                self.logits.data_ptr(), B, T, t, G, Ch, s)

    def stage_head(self):
        """(a8-a10) composed transposed-conv head: gaze_grcn.py:318-366."""
        if self.head != "deconv":
            return
        self._compose_head()
        taps = [(3 - dy) * HEAD_W + (3 - dx) for dy in range(4) for dx in range(4)]
        self._gemm(A=self.HS.data_ptr(), rowsA=self.HS.shape[0], lda=self.Ch, Bw=self.bhead.data_ptr(),
                   M=self.T * self.B * HEAD_S, N=48, K=self.Ch, taps=taps, epi=L.EPI_HEAD, S=HEAD_S, Wp=HEAD_W,
                   G=9, B=self.B, T=self.T, out_f32=self.logits.data_ptr(), bias=self._ptr("out_b"))

    def time_stages(self, c3d, gt, iters=3, emb=False):
        """CUDA-event time (ms, mean of `iters`) of each stage of one forward pass on the current stream."""
        names = ["embed", "i2h", "recurrence", "head", "loss"]
        fns = [(lambda: self.stage_embed(emb=c3d)) if emb else (lambda: self.stage_embed(c3d)), self.stage_i2h, self.stage_recurrence, self.stage_head,
               lambda: self.loss_fwd(gt)]
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

    def _compose_head(self):
        # under a graph capture of a TRAINING step the composition must be part of the graph (the parameters change
        # at every replay), whatever the host-side version counters say at capture time
        capturing = self.device.type == "cuda" and torch.cuda.is_current_stream_capturing()
        if self.head_version == self.param_version and not (capturing and self.train):
            return
        self.lib.rgp_head_compose(self._ptr("w1"), self._ptr("w2"), self._ptr("w3"), self._ptr("out_W"),
                                  self.w3f.data_ptr(), self.w23.data_ptr(), self.w123.data_ptr(),
                                  self.bhead.data_ptr(), self.Ch, self.dtype, self._s())
        self.head_version = self.param_version

    def loss_fwd(self, gt, loss_type="xentropy", global_batch=None):
        """global_batch: clips over all data-parallel ranks (the loss and its gradient are means over B_global*T)."""
        if loss_type not in LOSS_TYPES:
            raise NotImplementedError(str(loss_type))   # gaze_rnn.py:405-406
        assert tuple(gt.shape) == tuple(self.logits.shape) and gt.dtype == torch.float32
        n, hw = self.B * self.T, self.GH * self.GH
        self.lib.rgp_softmax_xent_fwd_bwd(self.logits.data_ptr(), gt.data_ptr(), self.probs.data_ptr(),
                                          self.dlogits.data_ptr() if self.train else 0,
                                          self.frame_loss.data_ptr(), n, hw, 1.0 / ((global_batch or self.B) * self.T),
                                          LOSS_TYPES[loss_type],
                                          self._s())
        self.lib.rgp_reduce_sum(self.frame_loss.data_ptr(), n, self.loss.data_ptr(), self._s())
        return self.loss

    # ------------------------------------------------------------------ backward (BPTT) — base.py:278-281
    def _wgrad(self, Y, ldy, ycol0, Ny, X, rowsX, ldx, Nx, M, taps, D):
        a = L.WgradArgs()
        a.Y, a.ldy, a.ycol0, a.Ny = Y, ldy, ycol0, Ny
        a.X, a.rowsX, a.ldx, a.xcol0, a.Nx = X, rowsX, ldx, 0, Nx
        a.M, a.ntaps = M, len(taps)
        for i, o in enumerate(taps):
            a.tap_off[i] = o
        a.D, a.dtype, a.engine = D, self.dtype, self.engine
        self.lib.rgp_wgrad(C.byref(a), self._s())

    def _mark(self, name):
        if self._prof is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._prof.append((name, ev))

    def time_backward(self, iters=3, **opt):
        """CUDA-event time (ms, mean of `iters`) of each section of backward() + optimizer_step()."""
        acc = OrderedDict()
        for _ in range(iters):
            self._prof = []
            self._mark("start")
            self.backward()
            self.optimizer_step(**opt)
            self._mark("clip_adam")
            torch.cuda.synchronize()
            for (n0, e0), (n1, e1) in zip(self._prof[:-1], self._prof[1:]):
                acc[n1] = acc.get(n1, 0.0) + e0.elapsed_time(e1) / iters
            self._prof = None
        return acc

    def backward(self):
        """Hand-written BPTT of the last forward(gt=...) (train=True): fills self.grads (flat, op layouts).
        dL/dlogits = (softmax - gt)/(B*T) was written by the fused softmax/loss kernel."""
        assert self.train, "construct the engine with train=True"
        lib, s = self.lib, self._s()
        B, T, S, Ch, Cin, Wp, G, P = self.B, self.T, self.S, self.Ch, self.Cin, self.Wp, self.G, self.P
        M = B * S
        es = self.dG.element_size()
        g = lambda name: self._ptr(name, self.grads)
        alpha = 1.0 / math.sqrt(1.0 + BN_EPS)
        lib.rgp_zero(self.grads.data_ptr(), self.grads.numel() * 4, s)
        # ---- head backward
        if self.head == "deconv":
            lib.rgp_head_patches(self.dlogits.data_ptr(), self.patches.data_ptr(), B, T, self.dtype, s)
            lib.rgp_head_w123t(self.w123.data_ptr(), self.w123T.data_ptr(), Ch, self.dtype, s)
            self._gemm(A=self.patches.data_ptr(), rowsA=self.patches.shape[0], lda=384, Bw=self.w123T.data_ptr(),
                       M=T * B * HEAD_S, N=Ch, K=384, taps=[0], epi=L.EPI_DHS, S=HEAD_S, Wp=HEAD_W, pado=HEAD_PAD,
                       G=G, So=S, Wo=Wp, out_f32=self.dhs.data_ptr(), ldo=Ch)
            lib.rgp_zero(self.dw123.data_ptr(), self.dw123.numel() * 4, s)
            self._wgrad(self.patches.data_ptr(), 384, 0, 384, self.HS.data_ptr(), self.HS.shape[0], Ch, Ch,
                        T * B * HEAD_S, [0], self.dw123.data_ptr())
            lib.rgp_head_compose_bwd(self.dw123.data_ptr(), self._ptr("w1"), self._ptr("w2"), self._ptr("w3"),
                                     self._ptr("out_W"), self.w3f.data_ptr(), self.w23.data_ptr(),
                                     self.dw23.data_ptr(), self.dw3f.data_ptr(), g("w1"), g("w2"), g("w3"),
                                     g("out_W"), Ch, s)
            lib.rgp_reduce_sum(self.dlogits.data_ptr(), self.dlogits.numel(), g("out_b"), s)
            if self.use_bn:
                lib.rgp_bn_bwd_all(self.dhs.data_ptr(), self._slab(self.Hf, 1), g("bn_gamma"), g("bn_beta"), alpha,
                                   B, T, G, Ch, s)
        else:
            if Ch % 4 == 0 and 256 % (Ch // 4) == 0:
                lib.rgp_fc_head_bwd_all(self.dlogits.data_ptr(), self._slab(self.Hf, 1), self._ptr("out_W"),
                                        self.dhs.data_ptr(), g("out_W"), g("out_b"), B, T, G, Ch, s)
            else:
                for t in range(T):
                    lib.rgp_fc_head_bwd(self.dlogits.data_ptr(), self._slab(self.Hf, t + 1), self._ptr("out_W"),
                                        self._slab(self.dhs, t), g("out_W"), g("out_b"), B, T, t, G, Ch, s)
        self._mark("head_bwd")
        btaps = [Wp + 1 - o for o in self.taps3]
        rowsT = T * M
        NG = self.ngx * Ch
        if self.lstm:
            self._backward_lstm_cell(btaps, g)
        else:
            self._backward_gru_cell(btaps, g, alpha)
        # ---- projection (only when the reference entry [B,T,1024,7,7] was used)
        if self.entry == "c3d":
            lib.rgp_transpose_taps(self._ptr("Wx_op"), self.WxT.data_ptr(), 9, NG, Cin, self.dtype, s)
            dense = (self.dtype == 1 and self.engine != 1 and os.environ.get("RGP_IM2COL", "1") != "0")
            self._gemm(A=self.dG.data_ptr(), rowsA=rowsT, lda=NG, K=NG, Bw=self.WxT.data_ptr(),
                       M=T * B * P if dense else rowsT, N=Cin, taps=btaps, epi=L.EPI_DX_DENSE, S=S, Wp=Wp, G=G, P=P,
                       out_op=self.dEmb.data_ptr(), ldo=Cin, im2col=2 if dense else 0)
            self._wgrad(self.dEmb.data_ptr(), Cin, 0, Cin, self.Apack.data_ptr(), self.Apack.shape[0], self.Cfeat,
                        self.Cfeat, T * B * P, [0], g("proj_W_op"))
            lib.rgp_colsum(self.dEmb.data_ptr(), Cin, T * B * P, Cin, g("proj_b"), self.dtype, s)
        self._mark("proj_bwd")

    def _backward_gru_cell(self, btaps, g, alpha):
        """BPTT through the ConvGRU steps (gaze_grcn.py:108-127) and the cell's weight gradients."""
        lib, s = self.lib, self._s()
        B, T, S, Ch, Cin, Wp, G = self.B, self.T, self.S, self.Ch, self.Cin, self.Wp, self.G
        M = B * S
        rowsT = T * M
        es = self.dG.element_size()
        # ---- transposed operand copies of the conv weights (dX = conv^T)
        lib.rgp_transpose_taps(self._ptr("Uc_op"), self.UcT.data_ptr(), 9, Ch, Ch, self.dtype, s)
        lib.rgp_transpose_taps(self._ptr("Uzr_op"), self.UzrT.data_ptr(), 9, 2 * Ch, Ch, self.dtype, s)
        self._mark("weight_transposes")
        if self.use_seq_small:
            # the whole reverse-time recurrence in ONE launch (rgp_convgru_seq_small_bwd: transposed weight slices resident in
            # shared memory, dL/dh's direct part in registers, two counter exchanges per step)
            lib.rgp_zero(self.seq_flags.data_ptr(), self.seq_flags.numel() * 4, s)
            a = L.GruBwdArgs()
            a.dG, a.rowsG = self.dG.data_ptr(), self.dG.shape[0]
            a.UcT, a.UzrT = self.UcT.data_ptr(), self.UzrT.data_ptr()
            a.Zs, a.Rs, a.Cs = self.Z.data_ptr(), self.Rs.data_ptr(), self.Cs.data_ptr()
            a.Hf, a.dhs = self.Hf.data_ptr(), self.dhs.data_ptr()
            a.bn_gamma = self._ptr("bn_gamma") if self.use_bn else 0
            a.alpha = alpha
            a.B, a.T, a.G, a.Ch, a.dtype = B, T, G, Ch, self.dtype
            lib.rgp_convgru_seq_small_bwd(C.byref(a), self.seq_flags.data_ptr(), s)
            self._mark("recurrence_bwd")
            self._wgrad_gru_cell(g)
            return
        # ---- recurrence, backwards in time: two tap GEMMs per step (mirror of the forward)
        gam = (lambda t: self._ptr("bn_gamma") + t * Ch * 4) if self.use_bn else (lambda t: 0)
        lib.rgp_gru_bwd_last(self._slab(self.dhs, T - 1), gam(T - 1), alpha, self.Z[T - 1].data_ptr(),
                             self.Cs[T - 1].data_ptr(), self._slab(self.Hf, T - 1), self._slab(self.dG, T - 1),
                             3 * Ch, self.dhpart.data_ptr(), B, G, Ch, self.dtype, s)
        for t in range(T - 1, -1, -1):
            dg_t = self._slab(self.dG, t)
            common = dict(M=M, rowsA=M, lda=3 * Ch, N=Ch, taps=btaps, S=S, Wp=Wp, G=G, Ch=Ch,
                          dhpart=self.dhpart.data_ptr())
            self._gemm(A=dg_t + 2 * Ch * es, K=Ch, Bw=self.UcT.data_ptr(), epi=L.EPI_BWD_RH,
                       save_r=self.Rs[t].data_ptr(), hprev_f32=self._slab(self.Hf, t), dg=dg_t, ldg=3 * Ch, **common)
            if t > 0:
                extra = dict(bias=gam(t - 1), alpha=alpha) if self.use_bn else {}
                self._gemm(A=dg_t, K=2 * Ch, Bw=self.UzrT.data_ptr(), epi=L.EPI_BWD_H,
                           dhs=self._slab(self.dhs, t - 1), save_z=self.Z[t - 1].data_ptr(),
                           save_c=self.Cs[t - 1].data_ptr(), hprev_f32=self._slab(self.Hf, t - 1),
                           dg=self._slab(self.dG, t - 1), ldg=3 * Ch, **common, **extra)
        self._mark("recurrence_bwd")
        self._wgrad_gru_cell(g)

    def _wgrad_gru_cell(self, g):
        """weight gradients of the ConvGRU cell: contractions over all T*B*S rows at once"""
        Ch, Cin = self.Ch, self.Cin
        rowsT = self.T * self.B * self.S
        self._wgrad(self.dG.data_ptr(), 3 * Ch, 0, 2 * Ch, self.Hop.data_ptr(), self.Hop.shape[0], Ch, Ch, rowsT,
                    self.taps3, g("Uzr_op"))
        self._wgrad(self.dG.data_ptr(), 3 * Ch, 2 * Ch, Ch, self.RH.data_ptr(), self.RH.shape[0], Ch, Ch, rowsT,
                    self.taps3, g("Uc_op"))
        self._wgrad(self.dG.data_ptr(), 3 * Ch, 0, 3 * Ch, self.Xp.data_ptr(), self.Xp.shape[0], Cin, Cin, rowsT,
                    self.taps3, g("Wx_op"))
        self._mark("wgrad_cell")

    def _backward_lstm_cell(self, btaps, g):
        """BPTT through the ConvLSTM steps as written in gaze_lstm.py:112-133 (W_hi shared by the input gate and the
        candidate, output gate peeping at the old cell, W_hc dead) and the cell's weight gradients."""
        lib, s = self.lib, self._s()
        B, T, S, Ch, Cin, Wp, G = self.B, self.T, self.S, self.Ch, self.Cin, self.Wp, self.G
        M = B * S
        rowsT = T * M
        lib.rgp_transpose_taps(self._ptr("Wh_op"), self.WhT.data_ptr(), 9, 3 * Ch, Ch, self.dtype, s)
        self._mark("weight_transposes")
        gfits = getattr(lib, "rgp_convgru_seq_small_fits", None)
        if self.use_seq_lstm and gfits is not None and gfits(B, G, Ch) and hasattr(lib, "rgp_convlstm_seq_small_bwd"):
            # the whole reverse-time recurrence in ONE launch (co-resident regime)
            lib.rgp_zero(self.seq_flags.data_ptr(), self.seq_flags.numel() * 4, s)
            a = L.LstmBwdArgs()
            a.dHc, a.rowsHc, a.dG4, a.WhT = self.dHc.data_ptr(), self.dHc.shape[0], self.dG.data_ptr(), self.WhT.data_ptr()
            a.save, a.Cf, a.dhs = self.Lsave.data_ptr(), self.Cf.data_ptr(), self.dhs.data_ptr()
            a.wci, a.wcf, a.wco = self._ptr("wci"), self._ptr("wcf"), self._ptr("wco")
            a.gwci, a.gwcf, a.gwco = g("wci"), g("wcf"), g("wco")
            a.B, a.T, a.G, a.Ch, a.dtype = B, T, G, Ch, self.dtype
            lib.rgp_convlstm_seq_small_bwd(C.byref(a), self.seq_flags.data_ptr(), s)
            self._mark("recurrence_bwd")
            self._wgrad_lstm_cell(g)
            return
        for t in range(T - 1, -1, -1):
            last = 1 if t == T - 1 else 0
            lib.rgp_lstm_gates_bwd(self.Lsave[t].data_ptr(), 0 if last else self.Lsave[t + 1].data_ptr(),
                                   self.Cf.data_ptr(), self._slab(self.dhs, t), self.dhrec.data_ptr(),
                                   self.dhpart.data_ptr(), self._ptr("wci"), self._ptr("wcf"), self._ptr("wco"),
                                   self._slab(self.dG, t), self._slab(self.dHc, t), g("wci"), g("wcf"), g("wco"),
                                   B, G, Ch, last, self.dtype, s)
            if t > 0:   # dL/dh_{t-1} through the three hidden convolutions
                self._gemm(A=self._slab(self.dHc, t), rowsA=M, lda=3 * Ch, K=3 * Ch, Bw=self.WhT.data_ptr(), M=M, N=Ch,
                           taps=btaps, epi=L.EPI_STORE_F32, out_f32=self.dhrec.data_ptr(), ldo=Ch, S=S, Wp=Wp, G=G)
        self._mark("recurrence_bwd")
        self._wgrad_lstm_cell(g)

    def _wgrad_lstm_cell(self, g):
        Ch, Cin = self.Ch, self.Cin
        rowsT = self.T * self.B * self.S
        self._wgrad(self.dHc.data_ptr(), 3 * Ch, 0, 3 * Ch, self.Hop.data_ptr(), self.Hop.shape[0], Ch, Ch, rowsT,
                    self.taps3, g("Wh_op"))
        self._wgrad(self.dG.data_ptr(), 4 * Ch, 0, 4 * Ch, self.Xp.data_ptr(), self.Xp.shape[0], Cin, Cin, rowsT,
                    self.taps3, g("Wx_op"))
        self._mark("wgrad_cell")

