"""GPU (-m gpu), collected last on purpose: b200_dcepp_forward_static -- the Zero-DCE++ forward whose scratch is a
file-scope `static __device__` array inside the library (no workspace argument, no allocation) -- called through the
C-ABI and checked against the existing entry point b200_dcepp_forward on the same inputs.  The two run the same
kernels on the same data, so the results are compared bit for bit (tighter than the 2.5e-4 the Zero-DCE++ parity
tests allow), on calls in a row that share the array with different shapes in between."""
import pytest
import torch

import etss_07_imageenhance_b200 as P
from etss_07_imageenhance_b200 import _lib
from oracle import oracle as O

pytestmark = pytest.mark.gpu
SF, ITERS = 12, 8


def _forward_pair(L, packed, x, with_adjust=True):
    """(enhanced, adjust) from the existing entry point and from the static-scratch one."""
    B, _, H, W = x.shape
    stream = torch.cuda.current_stream().cuda_stream
    wsb = L.b200_dcepp_workspace_bytes(B, H, W, SF)
    ws = torch.empty(wsb, dtype=torch.uint8, device=x.device)
    e0, a0 = torch.empty_like(x), torch.empty_like(x)
    e1, a1 = torch.full_like(x, -1.0), torch.full_like(x, -1.0)
    _lib.check(L.b200_dcepp_forward(x.data_ptr(), packed.data_ptr(), e0.data_ptr(), a0.data_ptr() if with_adjust else None,
                                    B, H, W, SF, ITERS, ws.data_ptr(), wsb, stream), "b200_dcepp_forward")
    _lib.check(L.b200_dcepp_forward_static(x.data_ptr(), packed.data_ptr(), e1.data_ptr(),
                                           a1.data_ptr() if with_adjust else None, B, H, W, SF, ITERS, stream),
               "b200_dcepp_forward_static")
    torch.cuda.synchronize()
    return (e0, a0), (e1, a1)


def test_dcepp_forward_static_matches_workspace_entry_point_on_calls_in_a_row(cuda_dev):
    L = _lib.lib()
    assert L.b200_dcepp_static_scratch_bytes() == 16 << 20
    m = P.ZeroDCEpp_RE(scale_factor=SF)
    m.load_state_dict(O.make_dcepp_state_dict(31, "stress"), strict=False)
    packed = m.to(cuda_dev).packed_weights()
    # calls in a row on the shared array: a 1080p frame, two small batches whose feature maps and flags land on the
    # first call's leftovers, then the first shape again
    shapes = [(1, 1080, 1920), (3, 36, 60), (2, 120, 252), (1, 1080, 1920)]
    for i, (B, H, W) in enumerate(shapes):
        assert L.b200_dcepp_workspace_bytes(B, H, W, SF) <= L.b200_dcepp_static_scratch_bytes()
        x = O.synthetic_lowlight(B, H, W, seed=900 + i).to(cuda_dev)
        (e0, a0), (e1, a1) = _forward_pair(L, packed, x)
        assert torch.equal(e1, e0), f"enhanced differs on call {i} {B, H, W}: {(e1 - e0).abs().max().item():.3e}"
        assert torch.equal(a1, a0), f"adjust differs on call {i} {B, H, W}: {(a1 - a0).abs().max().item():.3e}"
        assert (e1 != x).any()   # stress weights: the forward did something
    # without the adjust output, twice more back to back; the second of them must not take device memory
    x = O.synthetic_lowlight(1, 1080, 1920, seed=950).to(cuda_dev)
    (e0, _), (e1, _) = _forward_pair(L, packed, x, with_adjust=False)
    assert torch.equal(e1, e0)
    e2 = torch.empty_like(x)
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    _lib.check(L.b200_dcepp_forward_static(x.data_ptr(), packed.data_ptr(), e2.data_ptr(), None, 1, 1080, 1920, SF, ITERS,
                                           torch.cuda.current_stream().cuda_stream), "b200_dcepp_forward_static")
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert torch.equal(e2, e0)
    assert free_after >= free_before, (free_before, free_after)


def test_dcepp_forward_static_rejects_what_does_not_fit(cuda_dev):
    L = _lib.lib()
    x = torch.zeros(1, 3, 24, 24, device=cuda_dev)
    p = x.data_ptr()
    # one 4K frame at scale_factor 12 needs 35.8 MiB of scratch: a clear error, nothing launched
    assert L.b200_dcepp_workspace_bytes(1, 2160, 3840, SF) > L.b200_dcepp_static_scratch_bytes()
    rc = L.b200_dcepp_forward_static(p, p, p, None, 1, 2160, 3840, SF, ITERS, None)
    assert rc == _lib.ERR_TOO_LARGE and b"too large" in L.b200_strerror(rc)
    assert L.b200_dcepp_forward_static(p, p, p, None, 1, 50, 48, SF, ITERS, None) == _lib.ERR_BAD_SHAPE
    assert L.b200_dcepp_forward_static(None, p, p, None, 1, 48, 48, SF, ITERS, None) == _lib.ERR_NULL
    torch.cuda.synchronize()
