/* pcp_dev.h — development and test hooks of libpcp_b200.so.  NOT part of the drop-in boundary (include/pcp.h): nothing in
 * pointcloudprocessing_b200/ calls these on a product path; tests/ and scripts/ use them to pin single kernels (one tcgen05
 * MMA chain, one weight-gradient GEMM, the implicit pooled layer) in isolation and to time them.  Host pointers, synchronous. */
#ifndef PCP_DEV_H_
#define PCP_DEV_H_

#include "pcp.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Development aid: cycles block 0 / thread 0 of the last pcp_fps_f32 launch made under PCP_FPS_PROF=1 (25 points per
 * thread) spent in: distance loop, warp keys, __syncthreads, push, exchange wait, read-back (scripts/fps_phases.py). */
int pcp_fps_phase_cycles(unsigned long long out[8]);

/* Tensor-core self-test (used by tests only): D (128 x N) = A (128 x K) . B (N x K)^T through one
 * tcgen05.mma chain with bf16-rounded operands; validates the shared-memory operand image, the
 * descriptors and the TMEM read-back in isolation.  Host pointers, synchronous. */
int pcp_selftest_umma(const float *h_a, const float *h_b, int N, int K, float *h_d);
/* same with a CTA pair (cluster of 2, cta_group::2): D (256 x N) = A (256 x K) . B (N x K)^T */
int pcp_selftest_umma2(const float *h_a, const float *h_b, int N, int K, float *h_d);

/* weight-gradient self-test (tests only): dW (cin x cout) = A^T (cin x rows) . dZ (rows x cout) through the MN-major
 * tcgen05 split-K kernel of the bf16 training step; operands rounded to bf16, fp32 accumulate.  Host pointers. */
int pcp_selftest_wgrad(const float *h_a, const float *h_dz, int64_t rows, int cin, int cout, float *h_dw);
/* forward / input-gradient GEMM self-test (tests only): C (rows x n) (+)= A (rows x k) . W through the device-side
 * weight-image packer + tcgen05 GEMM of the bf16 training step; W is (k x n) row-major, or (n x k) row-major when
 * transpose_w (dA = dZ . W^T).  Host pointers; h_c is read when accumulate != 0. */
int pcp_selftest_tc_mm(const float *h_a, const float *h_w, int64_t rows, int k, int n, int transpose_w, int accumulate, float *h_c);
/* the same with split-K: the k / 64 k-tiles are dealt over ksplit CTAs that ADD their products into h_c with fp32 atomics
 * (the long, skinny input-gradient GEMMs of the T-Net tails) */
int pcp_selftest_tc_mm_splitk(const float *h_a, const float *h_w, int64_t rows, int k, int n, int transpose_w, int ksplit, float *h_c);
/* test hook: C[g] (m x n) = act(bf16(A[g]) . bf16(W) + bias[g]) with bf16 activations in HBM (large shapes run the persistent TMA
 * kernel, small ones the per-tile kernel); out_kind 0 fp32, 1 bf16, 2 softmax (n <= 32); h_c is fp32; h_bias (groups x n) or NULL */
int pcp_selftest_tc_gemm_bf16(const float *h_a, const float *h_w, const float *h_bias, int64_t m, int groups, int k, int n,
                              int out_kind, int relu, float *h_c);
/* development: average milliseconds of one generic tensor-core GEMM launch on device-resident synthetic operands
 * (out_kind 0 fp32, 1 bf16, 2 softmax; dbg 1 skips output stores, 2 skips activation loads) */
int pcp_bench_tc_gemm(int64_t rows, int k, int n, int out_kind, int a_f32, int relu, int iters, int dbg, float *h_ms);
/* max-pooled layer self-test (tests only): conv 128 -> cout, BN (batch statistics when batch != 0, else moving mean 0 /
 * variance 1), ReLU, reduce_max over the n points of each of the b clouds, and its backward pass, through the implicit
 * path of the bf16 training step (no rows x cout tensor; train_pool.inl).  Host pointers:
 * A (b*n x 128), W (128 x cout), gamma / beta (cout), dg (b x cout) -> g (b x cout), mean / inv (cout: batch mean and
 * 1/sqrt(var + 1e-3)), dW (128 x cout), dgamma / dbeta (cout), dA (b*n x 128). */
int pcp_selftest_pooled_layer(const float *h_a, const float *h_w, const float *h_gamma, const float *h_beta, const float *h_dg, int b,
                              int64_t n, int cout, int batch, float *h_g, float *h_mean, float *h_inv, float *h_dw, float *h_dgamma,
                              float *h_dbeta, float *h_da);
/* (tests only) the pcp_allreduce_fn callback (pcp.h) and row scale under which the following
 * pcp_selftest_pooled_layer calls run: emulates the data-parallel BatchNorm protocol of the pooled layer on one device. */
int pcp_selftest_set_bn_allreduce(int (*fn)(void *, void *, size_t, int, void *), void *user, double row_scale);

#ifdef __cplusplus
}
#endif
#endif /* PCP_DEV_H_ */
