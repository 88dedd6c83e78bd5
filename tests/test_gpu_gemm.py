"""Layer-wise tcgen05 GEMM (bfk_gemm) and operand packing against torch fp32 matmul on fp16-rounded operands.
Both operand roles, split-K with atomics, add-to-existing, bias, ragged M."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from bayesflow_b200 import _lib
    return _lib, _lib.lib()


def pack_rows(x, row_tile, k_pad=None):
    """x (R, K) fp32 cuda, contraction along columns"""
    L, lib = _lib()
    R, K = x.shape
    k_pad = k_pad or (K + 31) // 32 * 32
    out = torch.empty(lib.bfk_packed_operand_bytes(R, k_pad, row_tile), dtype=torch.uint8, device=x.device)
    L.check(lib.bfk_pack_kmajor(L.ptr(x), x.stride(0), R, K, k_pad, row_tile, 0, L.ptr(out), L.stream_ptr()), "pack")
    return out


def pack_cols(x, row_tile, k_pad=None):
    """x (K, F) fp32 cuda, contraction along rows; operand rows = columns of x"""
    L, lib = _lib()
    K, F = x.shape
    k_pad = k_pad or (K + 31) // 32 * 32
    out = torch.empty(lib.bfk_packed_operand_bytes(F, k_pad, row_tile), dtype=torch.uint8, device=x.device)
    L.check(lib.bfk_pack_kmajor_t(L.ptr(x), x.stride(0), K, k_pad, F, row_tile, 0, L.ptr(out), L.stream_ptr()), "pack_t")
    return out


def gemm(Ap, Bp, Cm, M, N, K, NT, splits=1, flags=0, bias=None, alpha=1.0):
    L, lib = _lib()
    L.check(lib.bfk_gemm(L.ptr(Ap), L.ptr(Bp), L.ptr(Cm), Cm.stride(0), M, N, K, NT, splits, flags, L.ptr(bias), alpha, 0,
                         L.stream_ptr()), "gemm")


def h(x):
    return x.half().float()


@pytest.mark.parametrize("M,K,N,NT", [(128, 256, 256, 256), (1000, 256, 256, 256), (333, 512, 256, 128), (4097, 1024, 1024, 256),
                                      (77, 64, 64, 64), (130, 96, 128, 128)])
def test_forward_role(M, K, N, NT):
    """Y = X W + b : A = X packed by rows, B = W (K,N) packed by columns"""
    g = torch.Generator(device="cuda").manual_seed(1)
    X = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(K, N, device="cuda", generator=g) / K ** 0.5
    b = torch.randn(N, device="cuda", generator=g)
    Y = torch.full((M, N), float("nan"), device="cuda")
    gemm(pack_rows(X, 128), pack_cols(W, NT), Y, M, N, (K + 31) // 32 * 32, NT, bias=b)
    ref = h(X) @ h(W) + b
    torch.testing.assert_close(Y, ref, atol=2e-4, rtol=2e-4)


def test_dx_role_and_add_existing():
    """dX = dY W^T + existing : A = dY packed by rows (contract over N), B rows = K index of W -> pack_rows(W, NT)"""
    M, K, N = 700, 256, 512
    g = torch.Generator(device="cuda").manual_seed(2)
    dY = torch.randn(M, N, device="cuda", generator=g)
    W = torch.randn(K, N, device="cuda", generator=g) / N ** 0.5
    base = torch.randn(M, K, device="cuda", generator=g)
    out = base.clone()
    gemm(pack_rows(dY, 128), pack_rows(W, 256), out, M, K, N, 256, flags=2)
    torch.testing.assert_close(out, base + h(dY) @ h(W).T, atol=2e-4, rtol=2e-4)


@pytest.mark.parametrize("splits", [1, 8, 37])
def test_dw_role_split_k(splits):
    """dW = X^T dY * alpha, contraction over the batch: A rows = features of X (pack_cols(X,128)), B rows = features of dY"""
    Bsz, K, N = 5000, 256, 256
    g = torch.Generator(device="cuda").manual_seed(3)
    X = torch.randn(Bsz, K, device="cuda", generator=g)
    dY = torch.randn(Bsz, N, device="cuda", generator=g)
    Bp = (Bsz + 31) // 32 * 32
    dW = torch.zeros(K, N, device="cuda")
    gemm(pack_cols(X, 128), pack_cols(dY, 256), dW, K, N, Bp, 256, splits=splits, flags=1 if splits > 1 else 0, alpha=1.0 / Bsz)
    ref = (h(X).T @ h(dY)) / Bsz
    torch.testing.assert_close(dW, ref, atol=2e-4, rtol=2e-3)


def test_bad_arguments_raise():
    L, lib = _lib()
    x = torch.zeros(128, 64, device="cuda")
    with pytest.raises(ValueError):
        gemm(pack_rows(x, 128), pack_rows(x, 64), torch.zeros(128, 64, device="cuda"), 128, 64, 64, 48)
