"""rgp_head_compose_static: the allocation-free variant of rgp_head_compose (scratch in a file-scope device array of the
library; no counterpart in the reference - include/rgp.h) against the existing entry point, through the C-ABI, on two
calls in a row.  Kept in a file pytest collects last."""
import numpy as np
import pytest
import torch

from recurrent_gaze_prediction_b200 import lib as L

pytestmark = pytest.mark.gpu


def _weights(rng, Ch):
    mk = lambda *s: torch.from_numpy(rng.uniform(-0.1, 0.1, size=s).astype(np.float32)).cuda()
    return mk(5, 5, 64, Ch), mk(5, 5, 32, 64), mk(7, 7, 12, 32), mk(12)


@pytest.mark.parametrize("dtype", [1, 0])
def test_head_compose_static_equals_head_compose_twice(dtype):
    lib = L.get_lib()
    s = torch.cuda.current_stream().cuda_stream
    tdt = torch.bfloat16 if dtype == 1 else torch.float32
    rng = np.random.RandomState(11)
    for Ch in (256, 128):           # two calls in a row on the shared scratch, the second with other shapes and weights
        w1, w2, w3, ow = _weights(rng, Ch)
        w3f, w23, w123 = (torch.zeros(n, device="cuda") for n in (49 * 32, 121 * 64, 361 * Ch))
        ref = torch.zeros(16 * 48, Ch, dtype=tdt, device="cuda")
        out = torch.full((16 * 48, Ch), 7.0, dtype=tdt, device="cuda")
        lib.rgp_head_compose(w1.data_ptr(), w2.data_ptr(), w3.data_ptr(), ow.data_ptr(), w3f.data_ptr(), w23.data_ptr(),
                             w123.data_ptr(), ref.data_ptr(), Ch, dtype, s)
        lib.rgp_head_compose_static(w1.data_ptr(), w2.data_ptr(), w3.data_ptr(), ow.data_ptr(), out.data_ptr(), Ch, dtype, s)
        torch.cuda.synchronize()
        assert float(ref.float().abs().max()) > 0
        assert torch.equal(out, ref)            # same kernels, same order: bit for bit


def test_head_compose_static_rejects_sizes_beyond_its_scratch():
    lib = L.get_lib()
    x = torch.zeros(16, device="cuda")
    with pytest.raises(Exception) as e:
        lib.rgp_head_compose_static(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), 1024, 1,
                                    torch.cuda.current_stream().cuda_stream)
    assert "512" in str(e.value)
