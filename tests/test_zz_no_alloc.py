"""The allocation-free entry point psgd_q_from_mt_noalloc (collected last on purpose: a run that stops at
the first failure has then already executed every other test).

It serves callers that cannot ask for device memory when they need Q = M^T P -- the GPU already belongs to
a framework's allocator, or the call sits in a loop that must not pay cudaMalloc's device-wide
synchronisation -- by keeping the split-n partial sums in a fixed array inside the library.  The test calls it
through the C ABI twice in a row, with different shapes so that the second call runs over whatever the first
left in the shared array, and compares each result with the planned entry point psgd_q_from_mt (workspace
allocated by the caller) within the tolerance of the other parity tests.  It also checks that nothing was
allocated through torch while the new entry point ran and that an oversized request is refused with a message.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-5  # same gate as tests/test_gpu_parity.py


def _planned(eng, M, P, r):
    """Q = M^T P through the existing entry point: a one-matrix plan and a caller-allocated workspace."""
    n, m = M.shape
    plan = eng.make_plan([(n, m)], r, [1])
    dummy = torch.zeros(n, m, device=M.device)
    v, vo = [torch.zeros(1, device=M.device)], [torch.zeros(1, device=M.device)]
    plan.set_ptrs([M], [dummy], [dummy.clone()], v, vo)
    q = torch.empty(m * r, device=M.device)
    plan.q_from_mt(P.reshape(-1).contiguous(), q)
    torch.cuda.synchronize()
    return q.view(m, r)


def _noalloc(lib, M, P, r):
    from network_distributed_pytorch_b200 import _lib

    n, m = M.shape
    q = torch.full((m, r), float("nan"), device=M.device)  # the caller passes the output
    stream = ctypes.c_void_p(torch.cuda.current_stream(M.device).cuda_stream)
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats(M.device)["allocation.all.allocated"]
    rc = lib.psgd_q_from_mt_noalloc(M.data_ptr(), P.data_ptr(), q.data_ptr(), n, m, r, M.device.index, stream)
    _lib.check(rc, lib)
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats(M.device)["allocation.all.allocated"] == before  # nothing came from torch either
    return q


def _rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def test_noalloc_matches_planned_entry_point_twice_in_a_row():
    from network_distributed_pytorch_b200 import _lib
    from network_distributed_pytorch_b200.engine import CudaEngine

    dev = torch.device("cuda", 0)
    eng = CudaEngine(dev)
    lib = _lib.load()
    gen = torch.Generator(device=dev).manual_seed(2024)
    # two calls in a row on different shapes: the second one reuses the array the first one filled
    for n, m, r in [(3000, 768, 4), (700, 1000, 8), (130, 147, 2)]:
        M = torch.randn(n, m, device=dev, generator=gen)
        P = torch.randn(n, r, device=dev, generator=gen)
        want = _planned(eng, M, P, r)
        got = _noalloc(lib, M, P, r)
        assert torch.isfinite(got).all()
        assert _rel(got, want) < TOL, (n, m, r, _rel(got, want))
        truth = M.double().t() @ P.double()
        assert _rel(got, truth) < TOL
        # and once more on the same input: the result does not depend on what the array held before
        again = _noalloc(lib, M, P, r)
        assert torch.equal(again, got)


def test_noalloc_refuses_what_does_not_fit():
    from network_distributed_pytorch_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda", 0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    t = torch.zeros(16, device=dev)
    # 2^20 rows x 768 columns at r = 4 would need 2048 row blocks x 768 x 4 x 4 bytes = 24 MiB of partials
    rc = lib.psgd_q_from_mt_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), 1 << 20, 768, 4, 0, stream)
    assert rc != 0
    msg = lib.psgd_last_error().decode()
    assert "scratch" in msg and "4194304" in msg
    assert lib.psgd_q_from_mt_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), 64, 64, 33, 0, stream) != 0
