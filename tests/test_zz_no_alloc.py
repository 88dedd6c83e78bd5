"""bfk_batch_moments_static (include/bfk.h): the allocation-free variant of bfk_batch_moments, its scratch in a file-scope
`static __device__` array.  No counterpart in the reference; the checker is
the existing entry point bfk_batch_moments on the same inputs.  Collected last on purpose (file name), so that under
`pytest -x` a failure here cannot hide another test."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, lib, x):
    rows, F = x.shape
    ws = torch.empty(int(L.bfk_moments_workspace_floats(rows, F)), device=x.device, dtype=torch.float32)
    want = torch.empty(2 * F + 1, device=x.device, dtype=torch.float32)
    got = torch.full((2 * F + 1,), float("nan"), device=x.device, dtype=torch.float32)
    lib.check(L.bfk_batch_moments(lib.ptr(x), rows, F, lib.ptr(ws), lib.ptr(want), lib.stream_ptr()), "bfk_batch_moments")
    lib.check(L.bfk_batch_moments_static(lib.ptr(x), rows, F, lib.ptr(got), lib.stream_ptr()), "bfk_batch_moments_static")
    torch.cuda.synchronize()
    return got, want


def test_static_scratch_entry_point_matches_the_workspace_one_twice_in_a_row():
    assert torch.cuda.is_available(), "the gpu-marked tests need a CUDA device"
    import bayesflow_b200._lib as lib
    L = lib.lib()
    assert int(L.bfk_static_scratch_floats()) == 1 << 19
    g = torch.Generator().manual_seed(3)
    # call 1: many blocks (fills a long prefix of the shared array); call 2: few rows, another F (stale contents behind it);
    # call 3: the first shape again with other data
    shapes = [(200_000, 32), (777, 10), (200_000, 32), (5, 3), (20_000, 257)]
    for rows, F in shapes:
        x = (torch.randn(rows, F, generator=g) * 3.0 + 1.5).cuda()
        got, want = _both(L, lib, x)
        assert torch.equal(got, want), (rows, F, float((got - want).abs().max()))  # same kernels, same order: bit-equal
        # and against torch (tolerance of tests/test_gpu_edge.py)
        torch.testing.assert_close(got[:F], x.mean(0), rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(got[F:2 * F], ((x - x.mean(0)) ** 2).sum(0), rtol=2e-4, atol=1e-3)
        assert float(got[2 * F]) == rows


def test_static_scratch_entry_point_allocates_nothing_and_rejects_oversize():
    import bayesflow_b200._lib as lib
    L = lib.lib()
    x = torch.randn(50_000, 16).cuda()
    out = torch.empty(33, device="cuda")
    lib.check(L.bfk_batch_moments_static(lib.ptr(x), 50_000, 16, lib.ptr(out), lib.stream_ptr()), "warm")
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        lib.check(L.bfk_batch_moments_static(lib.ptr(x), 50_000, 16, lib.ptr(out), lib.stream_ptr()), "static")
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0  # no device allocation by the driver-level allocator either
    # bad arguments fail loudly, with a message
    assert L.bfk_batch_moments_static(lib.ptr(x), 50_000, 2048, lib.ptr(out), lib.stream_ptr()) != 0
    assert b"bfk_batch_moments_static" in L.bfk_last_error_string()
    assert L.bfk_batch_moments_static(lib.ptr(x), 10, 16, None, lib.stream_ptr()) != 0
    # beyond the static array: rejected before anything is launched (so the oversized `rows` claim is never dereferenced)
    big_rows, big_F = 1 << 30, 1024
    assert int(L.bfk_moments_workspace_floats(big_rows, big_F)) > int(L.bfk_static_scratch_floats())
    assert L.bfk_batch_moments_static(lib.ptr(x), big_rows, big_F, lib.ptr(out), lib.stream_ptr()) == -69
    msg = L.bfk_last_error_string()
    assert b"scratch floats" in msg and b"524288" in msg, msg
    torch.cuda.synchronize()
