"""GPU: the tcgen05 building blocks in isolation — shared-memory operand image (K-major SWIZZLE_128B),
smem/instruction descriptors, TMEM read-back — against a NumPy bf16-rounded matmul."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def bf16_round(a):
    return torch.from_numpy(a).to(torch.bfloat16).to(torch.float32).numpy()


@pytest.mark.parametrize("N,K", [(64, 64), (128, 64), (128, 128), (256, 128), (32, 256)])
def test_umma_gemm_matches_bf16_matmul(lib, N, K):
    from pointcloudprocessing_b200 import _ffi
    rng = np.random.default_rng(N * 1000 + K)
    a = rng.standard_normal((128, K)).astype(np.float32)
    b = rng.standard_normal((N, K)).astype(np.float32)
    d = np.zeros((128, N), np.float32)
    _ffi.check(lib.pcp_selftest_umma(_ffi.fptr(a), _ffi.fptr(b), N, K, _ffi.fptr(d)))
    ref = bf16_round(a).astype(np.float64) @ bf16_round(b).astype(np.float64).T
    assert np.abs(d - ref).max() <= 1e-3 * np.abs(ref).max()


@pytest.mark.parametrize("N,K", [(64, 64), (128, 64), (128, 128), (256, 128)])
def test_umma_pair_gemm_matches_bf16_matmul(lib, N, K):
    """cta_group::2: one MMA across a CTA pair (M = 256; A rows and B rows split across the two CTAs)."""
    from pointcloudprocessing_b200 import _ffi
    rng = np.random.default_rng(N * 1000 + K + 1)
    a = rng.standard_normal((256, K)).astype(np.float32)
    b = rng.standard_normal((N, K)).astype(np.float32)
    d = np.zeros((256, N), np.float32)
    _ffi.check(lib.pcp_selftest_umma2(_ffi.fptr(a), _ffi.fptr(b), N, K, _ffi.fptr(d)))
    ref = bf16_round(a).astype(np.float64) @ bf16_round(b).astype(np.float64).T
    assert np.abs(d - ref).max() <= 1e-3 * np.abs(ref).max()


def _bf16_round(a):
    import torch
    return torch.from_numpy(a).to(torch.bfloat16).to(torch.float32).numpy()


@pytest.mark.parametrize("rows,cin,cout", [(64, 128, 128), (200, 64, 64), (1000, 128, 1024), (4096, 64, 128), (333, 512, 256), (96, 1024, 512),
                                            (70000, 128, 1024), (50, 256, 4096), (129, 64, 512)])
def test_wgrad_mn_major_tensor_core(lib, rows, cin, cout):
    """dW = A^T . dZ on tcgen05 with MN-major SWIZZLE_128B operands (tc_train.cu): exact up to fp32 summation order once
    the operands are rounded to bf16 like the kernel rounds them."""
    import ctypes as C
    rng = np.random.default_rng(rows + cin)
    a = rng.standard_normal((rows, cin)).astype(np.float32)
    dz = rng.standard_normal((rows, cout)).astype(np.float32)
    dw = np.zeros((cin, cout), np.float32)
    fp = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
    st = lib.pcp_selftest_wgrad(fp(a), fp(dz), rows, cin, cout, fp(dw))
    assert st == 0, lib.pcp_last_error()
    ref = _bf16_round(a).astype(np.float64).T @ _bf16_round(dz).astype(np.float64)
    err = np.abs(dw - ref).max() / np.abs(ref).max()
    assert err < 2e-5, err
