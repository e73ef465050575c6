"""gsr_l1_loss_noalloc (include/gsr_b200.h): the L1 loss + gradient kernel on scratch memory that lives inside the
library (a static __device__ array) instead of scratch the caller allocates.  No counterpart in the reference; the
entry point exists for callers that cannot call cudaMalloc.  Checked through the C-ABI against gsr_l1_loss (the
existing entry point, itself checked against the oracle in test_gpu_training.py) on two calls in a row: the second
call finds the first call's contents in the shared array.

This file is named so that pytest collects it LAST.
"""
import pytest
import torch

import gpu_utils as U
from gaussian_splatting_lightning_b200 import _native as N

pytestmark = pytest.mark.gpu


def _existing(a, b):
    L = N.lib()
    scratch = torch.zeros(L.gsr_l1_loss_scratch_bytes(), dtype=torch.uint8, device=a.device)
    grad, loss = torch.empty_like(a), torch.empty(1, device=a.device)
    N.check(L.gsr_l1_loss(a.data_ptr(), b.data_ptr(), a.numel(), grad.data_ptr(), loss.data_ptr(), scratch.data_ptr(),
                          N.stream_ptr(a.device)), "gsr_l1_loss")
    torch.cuda.synchronize()
    return loss, grad


def _noalloc(a, b, grad, loss):
    """outputs are allocated by the caller BEFORE the call; the call itself must not allocate"""
    N.check(N.lib().gsr_l1_loss_noalloc(a.data_ptr(), b.data_ptr(), a.numel(), grad.data_ptr() if grad is not None else None,
                                        loss.data_ptr(), N.stream_ptr(a.device)), "gsr_l1_loss_noalloc")


@pytest.mark.parametrize("shape", [(3, 1080, 1920), (3, 37, 53), (1, 1, 5)])
def test_noalloc_l1_loss_equals_the_existing_entry_point_twice_in_a_row(shape):
    g = torch.Generator().manual_seed(sum(shape))
    pairs = []
    for _ in range(2):     # two DIFFERENT inputs: the second call must not see anything of the first
        a = torch.rand(shape, generator=g).to(U.DEV)
        b = torch.rand(shape, generator=g).to(U.DEV)
        b.view(-1)[::7] = a.view(-1)[::7]          # exact zeros of the difference: sign(0) = 0
        pairs.append((a, b))
    for a, b in pairs:
        want_loss, want_grad = _existing(a, b)
        grad = torch.full_like(a, float("nan"))
        loss = torch.full((1,), float("nan"), device=a.device)
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats(a.device).get("num_device_alloc", 0)
        _noalloc(a, b, grad, loss)
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(a.device).get("num_device_alloc", 0) == before
        # same kernel, same grid, same deterministic summation order: bit-identical (the tolerance of
        # test_gpu_training.py against the oracle is 1e-6 relative; equality is the tighter statement)
        assert torch.equal(loss, want_loss), (float(loss), float(want_loss))
        assert torch.equal(grad, want_grad)


def test_noalloc_l1_loss_without_gradient_and_on_a_side_stream():
    a = torch.rand(3, 200, 300, device=U.DEV)
    b = torch.rand(3, 200, 300, device=U.DEV)
    want_loss, _ = _existing(a, b)
    loss = torch.zeros(1, device=U.DEV)
    s = torch.cuda.Stream(device=U.DEV)
    s.wait_stream(torch.cuda.current_stream(U.DEV))
    with torch.cuda.stream(s):
        for _ in range(3):         # back to back on one stream: stream order serialises the shared scratch
            _noalloc(a, b, None, loss)
    s.synchronize()
    assert torch.equal(loss, want_loss)


def test_noalloc_l1_loss_rejects_what_it_cannot_take():
    L = N.lib()
    a = torch.rand(16, device=U.DEV)
    loss = torch.zeros(1, device=U.DEV)
    too_many = (1 << 28) + 4
    rc = L.gsr_l1_loss_noalloc(a.data_ptr(), a.data_ptr(), too_many, None, loss.data_ptr(), N.stream_ptr(a.device))
    assert rc == -4                                                  # GSR_ELIMIT
    assert b"GSR_L1_NOALLOC_MAX_ELEMENTS" in L.gsr_last_error()
    assert L.gsr_l1_loss_noalloc(a.data_ptr(), a.data_ptr(), 0, None, loss.data_ptr(), N.stream_ptr(a.device)) == -1
    assert L.gsr_l1_loss_noalloc(a.data_ptr() + 4, a.data_ptr(), 8, None, loss.data_ptr(), N.stream_ptr(a.device)) == -1
    assert L.gsr_l1_loss_noalloc(None, a.data_ptr(), 8, None, loss.data_ptr(), N.stream_ptr(a.device)) == -1
