// gemm_simt.cu — fp32 CUDA-core GEMM, 128x64x16 tiles, 8x4 outputs per thread, fused
// bias / ReLU / per-cloud column max (the reduce_max of pointnet/PointNet.py:248,429 fused into the
// producing layer so the b x n x 1024 activation never exists).
#include "gemm_simt.cuh"

namespace pcp {

constexpr int BM = 128, BN = 64, BK = 16, GT = 256;

template <bool TA, bool TB>
__global__ void __launch_bounds__(GT)
gemm_f32_kernel(GemmArgs a)
{
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    __shared__ float smax[16][BN];

    const int tid = threadIdx.x;
    const int g = blockIdx.z / a.splitk, sk = blockIdx.z % a.splitk;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const float *A = a.A + (int64_t)g * a.strideA;
    const float *B = a.B + (int64_t)g * a.strideB;

    int kc = (a.K + a.splitk - 1) / a.splitk;
    kc = (kc + BK - 1) / BK * BK;
    const int kbeg = sk * kc, kend = min(a.K, kbeg + kc);

    const int tx = tid & 15, ty = tid >> 4;
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    const bool a_vec = (a.lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(A) & 15u) == 0);
    const bool b_vec = (a.ldb % 4 == 0) && ((reinterpret_cast<uintptr_t>(B) & 15u) == 0);

    for (int k0 = kbeg; k0 < kend; k0 += BK) {
        // ---- A tile ----
        if (!TA) {
            const int row = tid >> 1, kb = (tid & 1) * 8;
            const int m = m0 + row;
            const float *src = A + (int64_t)m * a.lda + k0 + kb;
            float v[8];
            if (m < a.M && a_vec && k0 + kb + 7 < kend) {
                const float4 q0 = __ldg(reinterpret_cast<const float4 *>(src));
                const float4 q1 = __ldg(reinterpret_cast<const float4 *>(src) + 1);
                v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w; v[4] = q1.x; v[5] = q1.y; v[6] = q1.z; v[7] = q1.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = (m < a.M && k0 + kb + j < kend) ? __ldg(src + j) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) As[kb + j][row] = v[j];
        } else {
            const int k = tid >> 4, mb = (tid & 15) * 8;
            const float *src = A + (int64_t)(k0 + k) * a.lda + m0 + mb;
            float v[8];
            if (k0 + k < kend && a_vec && m0 + mb + 7 < a.M) {
                const float4 q0 = __ldg(reinterpret_cast<const float4 *>(src));
                const float4 q1 = __ldg(reinterpret_cast<const float4 *>(src) + 1);
                v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w; v[4] = q1.x; v[5] = q1.y; v[6] = q1.z; v[7] = q1.w;
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = (k0 + k < kend && m0 + mb + j < a.M) ? __ldg(src + j) : 0.f;
            }
            *reinterpret_cast<float4 *>(&As[k][mb]) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4 *>(&As[k][mb + 4]) = make_float4(v[4], v[5], v[6], v[7]);
        }
        // ---- B tile ----
        if (!TB) {
            const int k = tid >> 4, nb = (tid & 15) * 4;
            const float *src = B + (int64_t)(k0 + k) * a.ldb + n0 + nb;
            float v[4];
            if (k0 + k < kend && b_vec && n0 + nb + 3 < a.N) {
                const float4 q = __ldg(reinterpret_cast<const float4 *>(src));
                v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = (k0 + k < kend && n0 + nb + j < a.N) ? __ldg(src + j) : 0.f;
            }
            *reinterpret_cast<float4 *>(&Bs[k][nb]) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            const int nn = tid >> 2, kb = (tid & 3) * 4;
            const float *src = B + (int64_t)(n0 + nn) * a.ldb + k0 + kb;
            float v[4];
            if (n0 + nn < a.N && b_vec && k0 + kb + 3 < kend) {
                const float4 q = __ldg(reinterpret_cast<const float4 *>(src));
                v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = (n0 + nn < a.N && k0 + kb + j < kend) ? __ldg(src + j) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) Bs[kb + j][nn] = v[j];
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[kk][ty * 8]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[kk][ty * 8 + 4]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[kk][tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[4] = {b0.x, b0.y, b0.z, b0.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }

    // ---- epilogue ----
    const float *bias = a.bias ? a.bias + (int64_t)g * a.strideBias : nullptr;
    float *C = a.C ? a.C + (int64_t)g * a.strideC : nullptr;
    float cmax[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int n = n0 + tx * 4 + j;
        const float bj = (bias && n < a.N) ? __ldg(bias + n) : 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int m = m0 + ty * 8 + i;
            float v = acc[i][j] * a.alpha + bj;
            if (a.relu) v = v > 0.f ? v : 0.f;
            if (m < a.M && n < a.N) {
                if (C) {
                    float *dst = C + (int64_t)m * a.ldc + n;
                    if (a.splitk > 1) atomicAdd(dst, v);
                    else if (a.accumulate) *dst += v;
                    else *dst = v;
                }
                cmax[j] = fmaxf(cmax[j], v);
            }
        }
    }
    if (a.colmax) {
#pragma unroll
        for (int j = 0; j < 4; ++j) smax[ty][tx * 4 + j] = cmax[j];
        __syncthreads();
        if (tid < BN) {
            float m = 0.f;
#pragma unroll
            for (int r = 0; r < 16; ++r) m = fmaxf(m, smax[r][tid]);
            const int n = n0 + tid;
            // post-ReLU values are >= 0, so the unsigned bit pattern orders like the float
            if (n < a.N) atomicMax(reinterpret_cast<unsigned int *>(a.colmax + (int64_t)g * a.strideColmax + n), __float_as_uint(m));
        }
    }
}

int gemm_f32(const GemmArgs &a, cudaStream_t st)
{
    PCP_REQUIRE(a.A && a.B && (a.C || a.colmax), PCP_ERR_INVALID, "gemm_f32: null operand");
    PCP_REQUIRE(!a.colmax || a.relu, PCP_ERR_INVALID, "gemm_f32: colmax needs relu");
    PCP_REQUIRE(a.splitk == 1 || (!a.bias && !a.relu && !a.colmax && !a.accumulate), PCP_ERR_INVALID, "gemm_f32: splitk with epilogue");
    if (a.M <= 0 || a.N <= 0 || a.groups <= 0) return PCP_OK;
    const int64_t gz = (int64_t)a.groups * a.splitk;
    PCP_REQUIRE(gz <= 65535, PCP_ERR_UNSUPPORTED, "gemm_f32: groups*splitk=%lld > 65535", (long long)gz);
    const int64_t gy = (a.M + BM - 1) / BM;
    PCP_REQUIRE(gy <= 65535, PCP_ERR_UNSUPPORTED, "gemm_f32: M=%d too large for one launch", a.M);
    dim3 grid((a.N + BN - 1) / BN, (unsigned)gy, (unsigned)gz);
    if (!a.transA && !a.transB) gemm_f32_kernel<false, false><<<grid, GT, 0, st>>>(a);
    else if (a.transA && !a.transB) gemm_f32_kernel<true, false><<<grid, GT, 0, st>>>(a);
    else if (!a.transA && a.transB) gemm_f32_kernel<false, true><<<grid, GT, 0, st>>>(a);
    else gemm_f32_kernel<true, true><<<grid, GT, 0, st>>>(a);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

}  // namespace pcp
