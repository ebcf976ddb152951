// tc_gemm.cuh — generic bf16 tensor-core GEMM (tcgen05 / TMEM) with the epilogues the PointNet heads need:
//   C[g] (M x N) = act( A[g] (M x K) . W[g] (K x N) + bias[g] )      A row-major (fp32 or bf16), W a packed image
// Used for the small-M dense layers (T-Net tails, classification head: pointnet/PointNet.py:431-442, :252-263)
// and the per-point segmentation head (:268-290) in the bf16 precision mode.
#pragma once
#include <vector>

#include "common.cuh"

namespace pcp {

struct TcGemmParams {
    const void *A = nullptr;            // (groups * M) x K row-major; fp32 if a_f32 else bf16
    int a_f32 = 0;
    int a_split3 = 0;                   // fp32 contract on bf16 tensor cores: K = 3 * Ksrc, the k-tiles are [hi(A) | hi(A) | lo(A)] of the
                                        // SAME fp32 A (lda columns Ksrc), against a weight image of [hi(W); lo(W); hi(W)] (pack_gemm_b_image_split3):
                                        // A.W ~ hi.hi + hi.lo + lo.hi, 16 mantissa bits instead of 8 (a_f32 must be set)
    int64_t lda = 0;                    // elements
    int64_t a_group_stride = 0;         // elements between groups
    const uint8_t *Bimg = nullptr;      // weight image from pack_gemm_b_image: tiles [nt][kt] of (BNt rows x 64 K) bf16 SW128
    int64_t b_group_stride = 0;         // bytes between per-group images (0 = shared weights)
    const float *bias = nullptr;        // per output column, may be null
    int64_t bias_group_stride = 0;      // floats (0 = shared)
    float *C32 = nullptr;               // optional fp32 output (row-major, ldc)
    __nv_bfloat16 *C16 = nullptr;       // optional bf16 output (row-major, ldc) — feeds the next layer
    int64_t ldc = 0;
    int64_t c_group_stride = 0;         // elements
    float *Csoftmax = nullptr;          // optional: softmax over the N (<= 32) columns, fp32 (row-major, ldc)
    int M = 0, N = 0, K = 0;            // M rows per group; K multiple of 64
    int groups = 1;
    int relu = 0;
    int accumulate = 0;                 // C32 += result (input-gradient accumulation); no bias / relu / softmax
    int ksplit = 1;                     // > 1: the K / 64 k-tiles are dealt over ksplit CTAs (grid.z) that ADD their products into C32 with
                                        // fp32 atomics — skinny, long-K products (a 256 x 4096 . 4096 x 256 gradient walked 64 dependent
                                        // k-tiles on 16 CTAs: 61 us); C32 only, no bias / relu; K / 64 must be a multiple of ksplit
    int dbg = 0;                        // development (pcp_bench_tc_gemm): 1 = skip output stores, 2 = skip activation loads
};

// order-preserving uint32 key of a float (larger float <=> larger key) and its inverse; packed extrema records of the
// training max-pool (tc_train.cu tc_pool_extrema): key << 32 | ~row, merged with a 64-bit atomicMax (first row wins ties)
__host__ __device__ inline uint32_t tc_float_key(uint32_t bits) { return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u); }
__host__ __device__ inline uint32_t tc_key_float_bits(uint32_t key) { return (key & 0x80000000u) ? (key ^ 0x80000000u) : ~key; }

// tile rows of the packed weight image for an N-column layer
inline int tc_gemm_bn_tile(int N) { return N >= 128 ? 128 : ((N + 15) / 16) * 16; }
// W: (K x N) row-major fp32 -> bf16 image (K % 64 == 0; N % 128 == 0 when N > 128)
void pack_gemm_b_image(const float *W, int K, int N, std::vector<uint16_t> &img);
size_t gemm_b_image_bytes(int K, int N);
// W: (K x N) row-major fp32 -> bf16 image of the (3K x N) matrix [hi(W); lo(W); hi(W)], hi = bf16(W), lo = bf16(W - hi)
void pack_gemm_b_image_split3(const float *W, int K, int N, std::vector<uint16_t> &img);

int tc_gemm(const TcGemmParams &p, cudaStream_t st);

}  // namespace pcp
