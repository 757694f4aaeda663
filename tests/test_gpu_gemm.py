"""GPU: the tcgen05/TMA GEMM mainloop in isolation (C-ABI opb200_debug_gemm) against fp64 matmul."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _planes(x):
    from onepose_st_b200 import _lib
    rows, cols = x.shape
    hi = torch.empty(rows, cols, dtype=torch.bfloat16, device=x.device)
    lo = torch.empty_like(hi)
    st = _lib.lib().opb200_split_planes(_lib.ptr(x), ctypes.c_int64(x.stride(0)), ctypes.c_int64(x.stride(1)),
                                        rows, cols, _lib.ptr(hi), _lib.ptr(lo), None, _lib.stream_ptr())
    _lib.check(st, "split_planes")
    return hi, lo


def test_split_planes_exact_and_transposed():
    x = torch.randn(300, 200, device="cuda")
    hi, lo = _planes(x)
    assert torch.equal(hi, x.to(torch.bfloat16))
    assert torch.equal(lo, (x - hi.float()).to(torch.bfloat16))
    xt = torch.randn(200, 300, device="cuda").t()  # column-major source
    hi2, lo2 = _planes(xt)
    assert torch.equal(hi2, xt.contiguous().to(torch.bfloat16))
    # hi + lo reproduces x to ~2^-17 relative
    assert ((hi.float() + lo.float()) - x).abs().max() <= x.abs().max() * 2 ** -16


@pytest.mark.parametrize("M,N,K,bn", [(128, 256, 64, 256), (128, 128, 256, 128), (1000, 768, 256, 256),
                                      (8064, 256, 512, 256), (333, 144, 128, 144), (4096, 512, 512, 128),
                                      (70000, 256, 256, 256)])
@pytest.mark.parametrize("prec", [1, 3])
def test_debug_gemm(M, N, K, bn, prec):
    from onepose_st_b200 import _lib
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g)
    b = torch.randn(N, K, device="cuda", generator=g)
    a_hi, a_lo = _planes(a)
    b_hi, b_lo = _planes(b)
    c = torch.full((M, N), float("nan"), device="cuda")
    st = _lib.lib().opb200_debug_gemm(_lib.ptr(a_hi), _lib.ptr(a_lo), _lib.ptr(b_hi), _lib.ptr(b_lo), M, N, K, bn, prec,
                                      _lib.ptr(c), _lib.stream_ptr())
    _lib.check(st, "debug_gemm")
    torch.cuda.synchronize()
    if prec == 1:
        ref = a_hi.double() @ b_hi.double().t()  # exact products of the bf16 values, fp32 accumulation
        tol = 1e-4 * K ** 0.5
    else:
        ref = a.double() @ b.double().t()
        tol = 5e-5 * K ** 0.5  # bf16x3 keeps ~2^-17 relative operand precision; sums are O(sqrt K), max over up to 18M outputs
    err = (c.double() - ref).abs().max().item()
    assert err < tol, f"max abs err {err} (tol {tol})"
