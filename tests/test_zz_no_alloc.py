"""GPU: the allocation-free OHEM entry point (rtseg_ohem_fwd_static) against rtseg_ohem_fwd.

Collected last on purpose (file name): a run that stops at the first failure has then already
exercised every other test.  The entry point serves callers that cannot allocate device memory —
a host in another runtime holding only its logits and labels, or a loop where cudaMalloc is
forbidden — by keeping its scratch in a device array inside the library.  This test calls it
straight through the C ABI with ctypes, twice in a row (the second call meets the first call's
leftovers in the shared array), and compares with the caller-allocated entry point.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, margin, shape):
    g = torch.Generator().manual_seed(seed)
    n, c, h, w = shape
    logits = torch.randn(shape, generator=g)
    labels = torch.randint(0, c, (n, h, w), generator=g)
    if margin:
        logits.scatter_add_(1, labels.unsqueeze(1), torch.full((n, 1, h, w), margin))
    labels[torch.rand((n, h, w), generator=g) < 0.1] = 255
    return logits.cuda(), labels.cuda()


def _reference_entry_point(lib, logits, labels, thresh):
    n, c, h, w = logits.shape
    loss_px = torch.empty(n * h * w, dtype=torch.float32, device="cuda")
    ws = torch.empty(lib.rtseg_ohem_workspace_bytes(), dtype=torch.uint8, device="cuda")
    out = torch.empty(4, dtype=torch.float32, device="cuda")
    rc = lib.rtseg_ohem_fwd(n, c, h, w, 0, 0, logits.data_ptr(), labels.data_ptr(), 255, thresh, loss_px.data_ptr(),
                            ws.data_ptr(), out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    return out.cpu()


def test_static_scratch_entry_point_matches_allocating_one_twice_in_a_row():
    from rtseg_b200 import _lib
    lib = _lib.load()
    _lib.sync_device(torch.cuda.current_device())
    thresh = float(-torch.log(torch.tensor(0.7)))
    assert lib.rtseg_ohem_fwd_static_max_pixels() == 1 << 21
    # call 1: big, "everything is hard" branch; call 2: smaller, top-k branch — it sees call 1's leftovers
    cases = [_case(1, 0.0, (2, 19, 96, 160)), _case(2, 12.0, (1, 19, 40, 56))]
    for logits, labels in cases:
        n, c, h, w = logits.shape
        out = torch.full((4,), -1.0, dtype=torch.float32, device="cuda")
        before = torch.cuda.memory_allocated()
        rc = lib.rtseg_ohem_fwd_static(n, c, h, w, 0, 0, logits.data_ptr(), labels.data_ptr(), 255, thresh,
                                       out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        assert rc == 0, _lib.last_error()
        assert torch.cuda.memory_allocated() == before  # nothing came from torch's allocator either
        got = out.cpu()
        want = _reference_entry_point(lib, logits, labels, thresh)
        # same kernels, same inputs: loss within the tolerance the other OHEM tests use, same branch and count
        assert abs(got[0].item() - want[0].item()) <= 1e-5 * max(1.0, abs(want[0].item()))
        assert got[2].item() == want[2].item() and got[3].item() == want[3].item()
        # and against the CPU oracle itself (reference core/loss.py:6-20 restated), not only against the sibling entry point
        from oracle import ref_ohem
        ref_loss = ref_ohem.ohem_ce_loss(logits.cpu(), labels.cpu(), 0.7)[0].item()
        assert abs(got[0].item() - ref_loss) <= 1e-5 * max(1.0, abs(ref_loss))
    assert {0.0, 1.0} == {float(_reference_entry_point(lib, lg, lb, thresh)[3]) for lg, lb in cases}


def test_static_scratch_entry_point_rejects_oversized_input():
    from rtseg_b200 import _lib
    lib = _lib.load()
    _lib.sync_device(torch.cuda.current_device())
    dummy = torch.zeros(8, device="cuda")
    out = torch.zeros(4, device="cuda")
    # 2 x 1025 x 1024 pixels > 2^21: must be refused before anything is launched
    rc = lib.rtseg_ohem_fwd_static(2, 19, 1025, 1024, 0, 0, dummy.data_ptr(), dummy.data_ptr(), 255, 0.3566,
                                   out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -2
    assert b"exceed" in lib.rtseg_last_error()
    assert ctypes.c_int64(lib.rtseg_ohem_fwd_static_max_pixels()).value == 2097152
