"""mt_loss_l1_static: the L1 loss entry point whose scratch is a file-scope device array (no allocation
inside the call; include/mt_b200.h).  It has no counterpart in the reference; it must agree with the existing
mt_loss_l1 (caller-provided scratch), which tests/kernel_cases.py checks against torch.

Collected last on purpose (file name)."""
import ctypes

import pytest

pytestmark = pytest.mark.gpu


def _l1_pair(L, a, b, weight, gscale):
    """(loss, da) from the existing entry point and from the no-allocation one, same inputs, same stream"""
    import torch
    n = a.numel()
    st = torch.cuda.current_stream().cuda_stream
    parts = L.mt_loss_l1_parts(n)
    scratch = torch.empty(parts, dtype=torch.float32, device=a.device)
    loss0 = torch.full((1,), float('nan'), device=a.device)
    loss1 = torch.full((1,), float('nan'), device=a.device)
    da0 = torch.full_like(a, float('nan'))
    da1 = torch.full_like(a, float('nan'))
    rc = L.mt_loss_l1(a.data_ptr(), b.data_ptr(), n, weight, gscale, loss0.data_ptr(), da0.data_ptr(),
                      scratch.data_ptr(), parts, st)
    assert rc == 0, L.mt_error_string(rc)
    rc = L.mt_loss_l1_static(a.data_ptr(), b.data_ptr(), n, weight, gscale, loss1.data_ptr(), da1.data_ptr(), st)
    assert rc == 0, L.mt_error_string(rc)
    torch.cuda.synchronize()
    return loss0, da0, loss1, da1


def test_l1_loss_with_built_in_scratch_matches_the_scratch_taking_entry_point():
    import torch
    from masterthesis_b200 import _lib
    L = _lib.lib()
    torch.manual_seed(0)
    dev = torch.device('cuda', 0)
    # two calls in a row: a large one (every one of the 296 partial slots in use), then a small ragged one
    # (3 blocks, n % 4 != 0) that finds the first call's sums still in the array
    for n, weight in ((16 * 3 * 256 * 256, 10.0), (2051, 0.5)):
        a = torch.randn(n, device=dev)
        b = torch.randn(n, device=dev)
        loss0, da0, loss1, da1 = _l1_pair(L, a, b, weight, weight)
        ref = weight * (a.double() - b.double()).abs().mean().item()
        # same kernels, same summation order: identical bits; and both within the op-level tolerance of
        # tests/test_kernels_gpu.py (1e-2 relative; measured ~1e-7) of a float64 evaluation
        assert torch.equal(loss0, loss1), (n, loss0.item(), loss1.item())
        assert torch.equal(da0, da1), n
        assert abs(loss1.item() - ref) <= 1e-2 * abs(ref), (n, loss1.item(), ref)
        assert torch.isfinite(da1).all()
    # without the optional gradient
    st = torch.cuda.current_stream().cuda_stream
    loss = torch.zeros(1, device=dev)
    rc = L.mt_loss_l1_static(a.data_ptr(), b.data_ptr(), a.numel(), 1.0, 1.0, loss.data_ptr(), None, st)
    assert rc == 0, L.mt_error_string(rc)
    torch.cuda.synchronize()
    assert abs(loss.item() - (a - b).abs().mean().item()) <= 1e-5


def test_l1_loss_with_built_in_scratch_rejects_sizes_beyond_its_limit():
    import torch
    from masterthesis_b200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda', 0)
    a = torch.zeros(16, device=dev)
    loss = torch.zeros(1, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    limit = L.mt_loss_l1_static_max_n()
    assert limit == 1 << 30
    rc = L.mt_loss_l1_static(a.data_ptr(), a.data_ptr(), limit + 1, 1.0, 1.0, loss.data_ptr(), None, st)
    assert rc == -1016
    msg = L.mt_error_string(rc)
    assert b"mt_loss_l1_static" in (msg if isinstance(msg, bytes) else msg.encode())
    rc = L.mt_loss_l1_static(a.data_ptr(), a.data_ptr(), 0, 1.0, 1.0, loss.data_ptr(), None, st)
    assert rc == -1005
    torch.cuda.synchronize()
