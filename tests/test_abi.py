"""The C-ABI library loads on a machine without a GPU and exports every symbol include/rgp.h declares."""
import ctypes
import os
import re

from recurrent_gaze_prediction_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "rgp.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rgp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    syms = header_symbols()
    assert len(syms) >= 25 and "rgp_tapgemm" in syms and "rgp_auc_judd" in syms
    dll = ctypes.CDLL(L.lib_path())
    missing = [s for s in syms if not hasattr(dll, s)]
    assert not missing, missing


def test_binding_covers_the_header():
    lib = L.RgpLib()
    assert sorted(lib.symbols()) == header_symbols()
    assert lib.version() >= 100


def test_struct_layouts_match_the_header():
    """sizeof of the ctypes mirrors == what the C compiler lays out (checked with gcc on the header)."""
    import subprocess
    import tempfile
    prog = '#include <stdio.h>\n#include "rgp.h"\nint main(){printf("%zu %zu %zu\\n", sizeof(rgp_tapgemm_args), sizeof(rgp_wgrad_args), sizeof(rgp_gru_seq_args));printf("%zu %zu %zu\\n", sizeof(rgp_lstm_seq_args), sizeof(rgp_gru_bwd_args), sizeof(rgp_lstm_bwd_args));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "s.c")
        open(c, "w").write(prog)
        exe = os.path.join(d, "s")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        a, b, c, d2, e2, f2 = map(int, subprocess.check_output([exe]).split())
    assert a == ctypes.sizeof(L.TapGemmArgs) and b == ctypes.sizeof(L.WgradArgs) and c == ctypes.sizeof(L.GruSeqArgs)
    assert d2 == ctypes.sizeof(L.LstmSeqArgs) and e2 == ctypes.sizeof(L.GruBwdArgs) and f2 == ctypes.sizeof(L.LstmBwdArgs)


def test_workspace_queries():
    lib = L.RgpLib()
    assert lib.rgp_convgru_seq_small_workspace_bytes(28, 7) == (2 * 14 + 1) * 4
    assert lib.rgp_convlstm_seq_workspace_bytes(8, 14) == (15 + 1) * 4
