// tc_train.cuh — tensor-core helpers of the bf16 training step (tc_train.cu)
#pragma once
#include "common.cuh"

namespace pcp {

// fp32 weights (row-major, leading dimension ldw) -> packed bf16 SWIZZLE_128B image of the logical (K x N) matrix
// B[k][n] = W[k*ldw + n] (transpose = false) or W[n*ldw + k] (transpose = true); layout of tc_gemm's Bimg
int pack_weight_image(const float *d_W, int K, int N, int ldw, bool transpose, uint8_t *d_img, cudaStream_t st);

// dW (cin x cout, leading dimension lddw) += A^T (cin x rows) . dZ (rows x cout): split-K over the rows, fp32 atomics
int tc_wgrad(const float *A, int64_t lda, int64_t rows, int cin, const float *dZ, int64_t ldz, int cout, float *dW, int64_t lddw,
             cudaStream_t st);

}  // namespace pcp
