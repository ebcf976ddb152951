// tail_fused.cuh — the three dense layers that follow a pooled feature (T-Net tails 1024->512->256->K^2, PointNet.py:431-442;
// classification head 1024->512->256->C + softmax, :252-263) as ONE launch for small batches (b <= 32: BASELINE config 1's batch).
#pragma once
#include <vector>

#include "common.cuh"

namespace pcp {

constexpr int kTailMaxRows = 32;

// W: (K x N) row-major fp32 -> bf16 in mma.sync B-fragment order (N padded to a multiple of 8, K % 32 == 0): for n-tile nt, k-pair kp,
// lane l one uint4 = { b0, b1 of k-step 2 kp ; b0, b1 of k-step 2 kp + 1 }
void pack_tail_weights(const float *W, int K, int N, std::vector<uint32_t> &out);

struct TailFusedParams {
    const float *A0 = nullptr;            // (b x 1024) fp32 pooled features
    int b = 0;
    const uint4 *w1 = nullptr, *w2 = nullptr, *w3 = nullptr;      // fragment-ordered weights of 1024->512, 512->256, 256->n3
    const float *t1 = nullptr, *t2 = nullptr, *t3 = nullptr;      // per-column bias (BatchNorm folded)
    int n3 = 0;                           // columns of the third layer
    float *out32 = nullptr;               // (b x n3) fp32, row-major, may be null
    __nv_bfloat16 *out16 = nullptr;       // (b x n3) bf16, row-major, may be null
    float *out_softmax = nullptr;         // (b x n3) softmax over the row (n3 <= 32), may be null
};

int tail_fused(const TailFusedParams &p, cudaStream_t st);

}  // namespace pcp
