// gemm_simt.cuh — fp32 CUDA-core GEMM with the fused epilogues the PointNet graph needs.
// This is the PCP_PRECISION_FP32 path (1e-3 contract) and the small-M / backward workhorse; the
// per-point layers of the bf16 path run in chain_tc.cu on tcgen05.
#pragma once
#include "common.cuh"

namespace pcp {

struct GemmArgs {
    // C[g] (M x N) = epi( op(A[g]) (M x K) . op(B[g]) (K x N) )
    const float *A = nullptr; int64_t lda = 0; int64_t strideA = 0;   // transA: element (m,k) at A[k*lda + m]
    const float *B = nullptr; int64_t ldb = 0; int64_t strideB = 0;   // transB: element (k,n) at B[n*ldb + k]
    float *C = nullptr; int64_t ldc = 0; int64_t strideC = 0;         // may be null (colmax-only)
    const float *bias = nullptr; int64_t strideBias = 0;              // per output column, per group stride
    float *colmax = nullptr; int64_t strideColmax = 0;                // per group (N) running max; needs relu, zero-init
    int M = 0, N = 0, K = 0;
    int groups = 1;
    int relu = 0;
    int transA = 0, transB = 0;
    int splitk = 1;            // >1: K split across blocks, atomicAdd into zero-initialised C (no bias/relu)
    float alpha = 1.0f;        // scales the product before bias
    int accumulate = 0;        // C += (non-atomic; splitk must be 1)
};

int gemm_f32(const GemmArgs &a, cudaStream_t st);

}  // namespace pcp
