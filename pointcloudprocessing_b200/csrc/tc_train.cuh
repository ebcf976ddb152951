// tc_train.cuh — tensor-core helpers of the bf16 training step (tc_train.cu)
#pragma once
#include "common.cuh"

namespace pcp {

// fp32 weights (row-major, leading dimension ldw) -> packed bf16 SWIZZLE_128B image of the logical (K x N) matrix
// B[k][n] = W[k*ldw + n] (transpose = false) or W[n*ldw + k] (transpose = true); layout of tc_gemm's Bimg.
// d_colscale (N floats, optional): B[k][n] is multiplied by colscale[n] before rounding (a frozen BatchNorm folded into its layer)
// d_sgn (N floats, optional): column n is multiplied by +1 (sgn[n] >= 0) or -1 (the pooled layers fold sign(gamma) into W)
int pack_weight_image(const float *d_W, int K, int N, int ldw, bool transpose, uint8_t *d_img, cudaStream_t st,
                      const float *d_colscale = nullptr, const float *d_sgn = nullptr);

// the same for a device table of jobs in ONE launch (no column scale); bnt = tc_gemm_bn_tile(N)
struct PackJob {
    const float *W; uint16_t *img;
    int K, N, ldw, transpose, bnt;
    const float *sgn;             // optional (N floats): column n is multiplied by +1 (sgn[n] >= 0) or -1 before rounding
};
int pack_weight_images(const PackJob *d_jobs, int njobs, cudaStream_t st);

// dW (cin x cout, leading dimension lddw) += A^T (cin x rows) . dZ (rows x cout): split-K over the rows, fp32 atomics
int tc_wgrad(const float *A, int64_t lda, int64_t rows, int cin, const float *dZ, int64_t ldz, int cout, float *dW, int64_t lddw,
             cudaStream_t st);
// deterministic Gram matrix G = A^T A (cin x cin) through per-split partial products summed in a fixed order
size_t tc_gram_scratch_floats(int cin);
int tc_gram_det(const float *A, int64_t lda, int64_t rows, int cin, float *G, float *scratch, size_t scratch_floats, cudaStream_t st);

// per-cloud column maxima of Z' = A . W' without storing Z': A (b*n x 128) fp32, Wimg = pack_weight_image(W, 128, cout, ..., sgn = gamma),
// i.e. W' = W . diag(sign(gamma)): relu(bn(.)) is monotone in sign(gamma) z, so the pooled row of a channel is the arg max of z'
// whatever the sign, and z = sign(gamma) z' exactly (one extremum to track instead of two: the epilogue was instruction bound).
// ext_max (b x cout) holds packed keys (tc_gemm.cuh: value key << 32 | ~row), zero-initialised by the caller
int tc_pool_extrema(const float *A, int b, int64_t n, const uint8_t *Wimg, int cout, unsigned long long *ext_max, cudaStream_t st);

// fp32-contract inference of a max-pooled 128 -> cout layer on the bf16 tensor cores (hi/lo split, see tc_train.cu):
// g (b x cout, zero-initialised) = max over each cloud's n rows of relu(A . W' + t); Whi / Wlo = pack_gemm_b_image of hi(W') / lo(W')
int tc_pool_max3(const float *A, int b, int64_t n, const uint8_t *Whi, const uint8_t *Wlo, const float *t, int cout, float *g, cudaStream_t st);

}  // namespace pcp
