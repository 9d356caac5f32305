#!/usr/bin/env python
"""Probe: UMMA K-major SWIZZLE_128B operand starting at a row that is not a multiple of the 8-row swizzle atom
(csrc/debug_umma.cu; build the library with RGP_BUILD_PROBES=1: the probe is not linked into the product .so).  Prints, per (row offset, base-offset mode), the max abs error against a CPU product."""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from recurrent_gaze_prediction_b200 import lib as L

dll = C.CDLL(L.lib_path())
fn = dll.rgp_debug_umma_rowoff
fn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
fn.restype = C.c_int
torch.manual_seed(0)
rows = 168
A = torch.randn(rows, 64, device="cuda").bfloat16()
B = torch.randn(128, 64, device="cuda").bfloat16()
D = torch.zeros(128, 128, device="cuda")
for off in (0, 8, 1, 3, 7, 9, 18, 33, 40):
    ref = A[off:off + 128].float() @ B.float().t()
    for mode in (0, 1):
        D.zero_()
        rc = fn(A.data_ptr(), rows, B.data_ptr(), D.data_ptr(), off, mode, None)
        torch.cuda.synchronize()
        err = (D - ref).abs().max().item()
        print("off=%2d mode=%d rc=%d max_err=%.4g %s" % (off, mode, rc, err, "OK" if err < 1e-2 else "WRONG"), flush=True)
