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


@pytest.mark.parametrize("rows,k,n,transpose,ksplit", [(256, 4096, 256, 1, 16), (100, 1024, 512, 1, 4), (300, 512, 64, 0, 8), (7, 1024, 9, 0, 2)])
def test_training_gemm_split_k(lib, rows, k, n, transpose, ksplit):
    """tc_gemm with the k-tiles dealt over `ksplit` CTAs that add their products with fp32 atomics (the T-Net tails' long,
    skinny input-gradient GEMMs: 4096 -> 256 over 256 rows walked 64 dependent k-tiles on 16 CTAs) — on top of what C holds."""
    import ctypes as C
    rng = np.random.default_rng(rows + k + n + ksplit)
    a = (rng.standard_normal((rows, k)) + 1.0).astype(np.float32)
    w = rng.standard_normal((n, k) if transpose else (k, n)).astype(np.float32)
    c0 = rng.standard_normal((rows, n)).astype(np.float32)
    c = c0.copy()
    fp = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
    st = lib.pcp_selftest_tc_mm_splitk(fp(a), fp(w), rows, k, n, transpose, ksplit, fp(c))
    assert st == 0, lib.pcp_last_error()
    wr = _bf16_round(w).astype(np.float64)
    ref = _bf16_round(a).astype(np.float64) @ (wr.T if transpose else wr) + c0
    err = np.abs(c - ref).max() / np.abs(ref).max()
    assert err < 2e-5, err


@pytest.mark.parametrize("rows,k,n,transpose,acc", [(768, 64, 64, 0, 0), (768, 64, 128, 0, 0), (768, 128, 1024, 0, 0), (6, 1024, 512, 0, 0),
                                                    (6, 256, 9, 0, 0), (6, 256, 4096, 0, 0), (768, 128, 8, 0, 0), (300, 512, 256, 0, 0),
                                                    (768, 64, 3, 1, 0), (768, 1024, 128, 1, 0), (768, 128, 64, 1, 1), (6, 512, 1024, 1, 0),
                                                    (1000, 256, 512, 1, 1), (100, 4096, 256, 1, 0)])
def test_training_gemm_device_packed_image(lib, rows, k, n, transpose, acc):
    """Forward (A . W) and input-gradient (dZ . W^T, optionally accumulating) GEMMs of the bf16 training step: weight image
    packed on the device every step (tc_train.cu pack_weight_image) + tc_gemm, against the float64 product of the
    bf16-rounded operands."""
    import ctypes as C
    rng = np.random.default_rng(rows * 7 + k + n)
    a = (rng.standard_normal((rows, k)) + 3.0).astype(np.float32)
    w = rng.standard_normal((n, k) if transpose else (k, n)).astype(np.float32)
    c0 = rng.standard_normal((rows, n)).astype(np.float32)
    c = c0.copy()
    fp = lambda x: x.ctypes.data_as(C.POINTER(C.c_float))
    st = lib.pcp_selftest_tc_mm(fp(a), fp(w), rows, k, n, transpose, acc, fp(c))
    assert st == 0, lib.pcp_last_error()
    wr = _bf16_round(w).astype(np.float64)
    ref = _bf16_round(a).astype(np.float64) @ (wr.T if transpose else wr) + (c0 if acc else 0.0)
    err = np.abs(c - ref).max() / np.abs(ref).max()
    assert err < 2e-5, err


@pytest.mark.parametrize("m,groups,k,n,kind,relu", [
    (1024, 40, 64, 512, 1, 1),        # seg_l1: per-cloud bias, 4 column tiles            (persistent TMA kernel: 1280 tiles)
    (40000, 1, 512, 256, 1, 1),       # ragged M (40000 = 312 * 128 + 64), 8 k-tiles, 2 column tiles
    (1000, 41, 128, 128, 0, 0),       # ragged per-group M: tiles straddle groups, rows past rows_valid must not be stored
    (50000, 1, 128, 12, 2, 0),        # soft-max epilogue, 16-row weight tile
    (50000, 1, 256, 72, 1, 1),        # N not a multiple of 32: slow store path
    (300, 3, 128, 128, 1, 1),         # small: per-tile kernel
])
def test_tc_gemm_bf16_activations(lib, m, groups, k, n, kind, relu):
    """The bf16-activation GEMM of the per-point layers against a float64 product of the bf16-rounded operands."""
    from pointcloudprocessing_b200 import _ffi
    rng = np.random.default_rng(m + k + n)
    a = rng.standard_normal((groups, m, k)).astype(np.float32)
    w = (rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)
    bias = rng.standard_normal((groups, n)).astype(np.float32)
    c = np.full((groups, m, n), np.nan, np.float32)
    _ffi.check(lib.pcp_selftest_tc_gemm_bf16(_ffi.fptr(a), _ffi.fptr(w), _ffi.fptr(bias), m, groups, k, n, kind, relu, _ffi.fptr(c)))
    r16 = lambda x: torch.from_numpy(x).bfloat16().double().numpy()
    ref = r16(a) @ r16(w) + bias[:, None, :].astype(np.float64)
    if relu:
        ref = np.maximum(ref, 0)
    if kind == 2:
        ex = np.exp(ref - ref.max(-1, keepdims=True))
        ref = ex / ex.sum(-1, keepdims=True)
    tol = 1e-2 if kind == 1 else 2e-4                       # bf16 output rounding vs fp32 accumulate
    assert np.isfinite(c).all()
    assert np.abs(c - ref).max() <= tol * max(1.0, np.abs(ref).max()), float(np.abs(c - ref).max())
