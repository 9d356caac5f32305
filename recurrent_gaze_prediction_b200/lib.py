"""ctypes binding of ``librgp_sm100.so`` (the C-ABI declared in ``include/rgp.h``).

The product path has no CPU fallback: if the shared library is missing or a call
fails, a ``RuntimeError`` is raised.  PyTorch is used by the callers only to own
device memory and streams; every pointer handed to the library is a raw device
address (``tensor.data_ptr()``).
"""
from __future__ import annotations

import ctypes as C
import os

RGP_MAX_TAPS = 32
(EPI_STORE_F32, EPI_PROJ, EPI_GRU_ZR, EPI_GRU_C, EPI_HEAD, EPI_STORE_OP_P, EPI_STORE_OP,
 EPI_BWD_RH, EPI_BWD_H, EPI_DHS, EPI_DX_DENSE, EPI_UP7, EPI_UP7X) = range(13)

_vp, _i, _ll, _f, _d = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_double


class TapGemmArgs(C.Structure):
    """Mirror of ``rgp_tapgemm_args`` (include/rgp.h) — field order is the ABI."""
    _fields_ = [
        ("A", _vp), ("rowsA", _ll), ("lda", _i),
        ("Bw", _vp),
        ("M", _i), ("N", _i), ("K", _i), ("ntaps", _i),
        ("tap_off", _i * RGP_MAX_TAPS),
        ("dtype", _i), ("engine", _i), ("epi", _i), ("accumulate", _i),
        ("S", _i), ("Wp", _i), ("G", _i), ("P", _i),
        ("So", _i), ("Wo", _i), ("pado", _i), ("Ch", _i),
        ("B", _i), ("T", _i), ("t", _i),
        ("out_f32", _vp), ("ldo", _i),
        ("out_op", _vp), ("out_op2", _vp),
        ("xg", _vp), ("ldx", _i),
        ("hprev_f32", _vp), ("hnext_f32", _vp), ("z_f32", _vp),
        ("save_r", _vp), ("save_c", _vp),
        ("bias", _vp), ("bias2", _vp), ("alpha", _f),
        ("dhpart", _vp), ("dhs", _vp), ("save_z", _vp), ("dg", _vp), ("ldg", _i),
        ("im2col", _i),
    ]


class WgradArgs(C.Structure):
    """Mirror of ``rgp_wgrad_args`` (include/rgp.h)."""
    _fields_ = [
        ("Y", _vp), ("ldy", _i), ("ycol0", _i), ("Ny", _i),
        ("X", _vp), ("rowsX", _ll), ("ldx", _i), ("xcol0", _i), ("Nx", _i),
        ("M", _i), ("ntaps", _i), ("tap_off", _i * RGP_MAX_TAPS),
        ("D", _vp),
        ("dtype", _i), ("engine", _i),
    ]


class GruSeqArgs(C.Structure):
    """Mirror of ``rgp_gru_seq_args`` (include/rgp.h)."""
    _fields_ = [
        ("Hop", _vp), ("rowsH", _ll), ("Hf", _vp),
        ("RH", _vp), ("rowsRH", _ll),
        ("Uzr", _vp), ("Uc", _vp), ("Xg", _vp),
        ("Zs", _vp), ("Rs", _vp), ("Cs", _vp),
        ("HS", _vp),
        ("bn_gamma", _vp), ("bn_beta", _vp), ("alpha", _f),
        ("B", _i), ("T", _i), ("G", _i), ("Ch", _i),
        ("So", _i), ("Wo", _i), ("pado", _i),
        ("nsplit", _i), ("dtype", _i), ("mcast", _i), ("ntile", _i),
    ]


class GruBwdArgs(C.Structure):
    """Mirror of ``rgp_gru_bwd_args`` (include/rgp.h)."""
    _fields_ = [
        ("dG", _vp), ("rowsG", _ll), ("UcT", _vp), ("UzrT", _vp),
        ("Zs", _vp), ("Rs", _vp), ("Cs", _vp), ("Hf", _vp), ("dhs", _vp),
        ("bn_gamma", _vp), ("alpha", _f),
        ("B", _i), ("T", _i), ("G", _i), ("Ch", _i), ("dtype", _i),
    ]


class LstmBwdArgs(C.Structure):
    """Mirror of ``rgp_lstm_bwd_args`` (include/rgp.h)."""
    _fields_ = [
        ("dHc", _vp), ("rowsHc", _ll), ("dG4", _vp), ("WhT", _vp), ("save", _vp), ("Cf", _vp), ("dhs", _vp),
        ("wci", _vp), ("wcf", _vp), ("wco", _vp), ("gwci", _vp), ("gwcf", _vp), ("gwco", _vp),
        ("B", _i), ("T", _i), ("G", _i), ("Ch", _i), ("dtype", _i),
    ]


class LstmSeqArgs(C.Structure):
    """Mirror of ``rgp_lstm_seq_args`` (include/rgp.h)."""
    _fields_ = [
        ("Hop", _vp), ("rowsH", _ll), ("Hf", _vp), ("Cf", _vp), ("Wh", _vp), ("Xg", _vp),
        ("wci", _vp), ("wcf", _vp), ("wco", _vp), ("HS", _vp), ("save", _vp), ("flags", _vp),
        ("B", _i), ("T", _i), ("G", _i), ("Ch", _i),
        ("So", _i), ("Wo", _i), ("pado", _i), ("dtype", _i),
    ]


_SIGS = {
    "rgp_tapgemm": [C.POINTER(TapGemmArgs), _vp],
    "rgp_convgru_seq_fwd": [C.POINTER(GruSeqArgs), _vp],
    "rgp_convgru_seq_profile": [_vp, _i],
    "rgp_convgru_seq_small_fwd": [C.POINTER(GruSeqArgs), _vp, _vp],
    "rgp_convgru_seq_small_bwd": [C.POINTER(GruBwdArgs), _vp, _vp],
    "rgp_convlstm_seq_fwd": [C.POINTER(LstmSeqArgs), _vp],
    "rgp_convlstm_seq_profile": [_vp, _i],
    "rgp_convlstm_seq_small_bwd": [C.POINTER(LstmBwdArgs), _vp, _vp],
    "rgp_pack_c3d": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "rgp_pack_c3d_src": [_vp, _i, _vp, _i, _i, _i, _i, _i, _vp],
    "rgp_pack_emb": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "rgp_cast": [_vp, _vp, _ll, _i, _vp],
    "rgp_fc_head": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "rgp_fc_head_all": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "rgp_lstm_gates": [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp],
    "rgp_head_compose": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
    "rgp_head_compose_static": [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp],
    "rgp_softmax_xent_fwd_bwd": [_vp, _vp, _vp, _vp, _vp, _i, _i, _f, _i, _vp],
    "rgp_reduce_sum": [_vp, _ll, _vp, _vp],
    "rgp_sumsq": [_vp, _ll, _vp, _vp],
    "rgp_gradnorm_clip_adam": [_vp, _vp, _vp, _vp, _ll, _vp, _f, _f, _f, _f, _f, _vp, _i, _vp],
    "rgp_adam_schedule": [_vp, _vp, _d, _i, _d, _d, _vp, _vp],
    "rgp_gradnorm_clip_adam_dev": [_vp, _vp, _vp, _vp, _ll, _vp, _f, _vp, _f, _f, _f, _vp, _i, _vp],
    "rgp_gradnorm_clip_momentum_dev": [_vp, _vp, _vp, _vp, _ll, _vp, _f, _vp, _i, _f, _f, _f, _vp, _i, _vp],
    "rgp_zero": [_vp, _ll, _vp],
    "rgp_flip_w": [_vp, _vp, _i, _ll, _i, _vp],
    "rgp_wgrad": [C.POINTER(WgradArgs), _vp],
    "rgp_gru_bwd_last": [_vp, _vp, _f, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _i, _vp],
    "rgp_lstm_gates_bwd": [_vp] * 14 + [_i, _i, _i, _i, _i, _vp],
    "rgp_transpose_taps": [_vp, _vp, _i, _i, _i, _i, _vp],
    "rgp_head_patches": [_vp, _vp, _i, _i, _i, _vp],
    "rgp_head_compose_bwd": [_vp] * 13 + [_i, _vp],
    "rgp_head_w123t": [_vp, _vp, _i, _i, _vp],
    "rgp_bn_bwd": [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _vp],
    "rgp_bn_bwd_all": [_vp, _vp, _vp, _vp, _f, _i, _i, _i, _i, _vp],
    "rgp_colsum": [_vp, _i, _i, _i, _vp, _i, _vp],
    "rgp_fc_head_bwd": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp],
    "rgp_fc_head_bwd_all": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    # dense-GRU variant (gaze_rnn.py:283-359)
    "rgp_dense_pack": [_vp, _i, _vp, _vp, _i, _ll, _i, _i, _i, _vp],
    "rgp_dense_pack_dropout": [_vp, _i, _vp, _vp, _i, _ll, _i, _i, _f, C.c_ulonglong, _i, _vp],
    "rgp_dense_gru_ru": [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "rgp_dense_gru_c": [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "rgp_dense_dlogits": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "rgp_dense_gru_bwd_c": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "rgp_dense_gru_bwd_r": [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp],
    "rgp_dense_unpack": [_vp, _i, _vp, _ll, _i, _i, _f, C.c_ulonglong, _i, _vp],
    "rgp_dense_logits": [_vp, _i, _vp, _vp, _i, _i, _i, _vp],
    # cascade top level (gaze_grcn_cascade.py:343-424)
    "rgp_topgru_seq_fwd": [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp, _i, _i, _i, _vp],
    "rgp_cascade_compose": [_vp, _i, _vp, _vp, _i, _i, _vp],
    "rgp_maxout": [_vp, _i, _vp, _i, _i, _vp, _i, _vp, _i, _i, _i, _vp],
    "rgp_fill_uniform_f64": [_vp, _ll, C.c_ulonglong, _vp],
    "rgp_fill_randint_rows": [_vp, _vp, _i, _i, _i, _i, C.c_ulonglong, _vp],
    # saliency metrics (evaluation_metrics.py)
    "rgp_count_nonzero_rows": [_vp, _i, _ll, _vp, _vp],
    "rgp_metrics_prep": [_vp, _vp, _i, _i, _i, _vp],
    "rgp_resize_bicubic": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "rgp_sim_cc_nss": [_vp, _vp, _vp, _vp, _i, _i, _vp],
    "rgp_auc_judd": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp],
    "rgp_auc_sampled": [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _d, _i, _i, _vp],
    "rgp_sim_cc_nss_resized": [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "rgp_auc_judd_resized": [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp],
    "rgp_auc_sampled_resized": [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _d, _i, _i, _vp],
}


def lib_path():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "librgp_sm100.so")


class RgpLib(object):
    """Loaded library; every C entry point is a method that raises on a non-zero return."""

    def __init__(self, path=None):
        path = path or lib_path()
        if not os.path.exists(path):
            raise RuntimeError("librgp_sm100.so not found at %s — run `python -c 'import __graft_entry__ as g; "
                               "g.build()'` (there is no CPU fallback)" % path)
        self._dll = C.CDLL(path)
        self._dll.rgp_last_error.restype = C.c_char_p
        self._dll.rgp_version.restype = _i
        self._dll.rgp_has_tcgen05.restype = _i
        self._dll.rgp_metrics_fused_fits.restype = _i
        self._dll.rgp_metrics_fused_fits.argtypes = [_i, _i, _i, _i]
        self._dll.rgp_convlstm_seq_fits.restype = _i
        self._dll.rgp_convlstm_seq_fits.argtypes = [_i, _i, _i]
        self._dll.rgp_convgru_seq_small_fits.restype = _i
        self._dll.rgp_convgru_seq_small_fits.argtypes = [_i, _i, _i]
        for nm in ("rgp_convgru_seq_small_workspace_bytes", "rgp_convlstm_seq_workspace_bytes"):
            getattr(self._dll, nm).restype = _ll
            getattr(self._dll, nm).argtypes = [_i, _i]
        for name, sig in _SIGS.items():
            fn = getattr(self._dll, name)
            fn.argtypes = sig
            fn.restype = _i
            setattr(self, name, self._wrap(name, fn))
        self.path = path
        self.launches = 0

    def _wrap(self, name, fn):
        def call(*args):
            rc = fn(*args)
            self.launches += 1
            if rc != 0:
                raise RuntimeError("%s failed (%d): %s" % (name, rc, self._dll.rgp_last_error().decode()))
            return rc
        return call

    def version(self):
        return self._dll.rgp_version()

    def has_tcgen05(self):
        return bool(self._dll.rgp_has_tcgen05())

    def rgp_metrics_fused_fits(self, h, w, oh, ow):
        """a predicate, not a status: 1 if the fused resize+metric kernels can hold an h x w -> oh x ow enlargement"""
        return int(self._dll.rgp_metrics_fused_fits(h, w, oh, ow))

    def rgp_convlstm_seq_fits(self, B, G, Ch):
        """a predicate: 1 if the persistent ConvLSTM kernel can hold this shape (all its CTAs co-resident)"""
        return int(self._dll.rgp_convlstm_seq_fits(B, G, Ch))

    def rgp_convgru_seq_small_fits(self, B, G, Ch):
        """a predicate: 1 if the small-batch persistent ConvGRU kernel can hold this shape"""
        return int(self._dll.rgp_convgru_seq_small_fits(B, G, Ch))

    def rgp_convgru_seq_small_workspace_bytes(self, B, G):
        return int(self._dll.rgp_convgru_seq_small_workspace_bytes(B, G))

    def rgp_convlstm_seq_workspace_bytes(self, B, G):
        return int(self._dll.rgp_convlstm_seq_workspace_bytes(B, G))

    def symbols(self):
        return sorted(list(_SIGS) + ["rgp_last_error", "rgp_version", "rgp_has_tcgen05", "rgp_metrics_fused_fits",
                                     "rgp_convlstm_seq_fits", "rgp_convgru_seq_small_fits", "rgp_convgru_seq_small_workspace_bytes",
                                     "rgp_convlstm_seq_workspace_bytes"])


_LIB = None


def get_lib():
    global _LIB
    if _LIB is None:
        _LIB = RgpLib()
    return _LIB
