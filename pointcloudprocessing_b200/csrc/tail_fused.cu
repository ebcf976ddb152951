// tail_fused.cu — small-batch dense tails in one launch (see tail_fused.cuh).
//
// At b = 32 (BASELINE config 1's batch) the three dense layers behind every pooled feature are latency, not work: as three
// dependent tc_gemm launches on 1 - 16 CTAs each they cost ~37 us, nine of the thirteen launches of a forward.  Here ONE cluster
// of 16 CTAs runs all three: every CTA owns 1/16 of the columns of each layer, streams its weight slice straight from L2 into
// registers in mma.sync B-fragment order (no shared-memory staging, 16-byte coalesced loads), keeps the activations of the whole
// (<= 32-row) batch in shared memory, and hands its slice of a layer's output to all peers through distributed shared memory
// before a cluster barrier.  bf16 operands, fp32 accumulation and bias, like tc_gemm.
#include <cooperative_groups.h>
#include <cuda_bf16.h>

#include <cmath>
#include <cstring>

#include "tail_fused.cuh"

namespace cg = cooperative_groups;

namespace pcp {

constexpr int kTailCluster = 16, kTailThreads = 256;
constexpr int kLd1 = 1024 + 8, kLd2 = 512 + 8, kLd3 = 256 + 8;      // bf16 row strides: +16 B so ldmatrix rows fall in distinct bank groups

static inline uint16_t bf16_bits_host(float f)
{
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);      // NaN
    u += 0x7fffu + ((u >> 16) & 1u);                                               // round to nearest even
    return (uint16_t)(u >> 16);
}

void pack_tail_weights(const float *W, int K, int N, std::vector<uint32_t> &out)
{
    const int NT = (N + 7) / 8, KP = K / 32;
    out.assign((size_t)NT * KP * 32 * 4, 0u);
    auto at = [&](int k, int n) -> uint32_t { return n < N ? bf16_bits_host(W[(size_t)k * N + n]) : 0u; };
    for (int nt = 0; nt < NT; ++nt)
        for (int kp = 0; kp < KP; ++kp)
            for (int lane = 0; lane < 32; ++lane) {
                const int n = nt * 8 + lane / 4, kq = (lane % 4) * 2;
                uint32_t *o = out.data() + ((size_t)(nt * KP + kp) * 32 + lane) * 4;
                for (int h = 0; h < 2; ++h) {            // the two k-steps of the pair
                    const int k0 = (2 * kp + h) * 16;
                    o[2 * h] = at(k0 + kq, n) | (at(k0 + kq + 1, n) << 16);               // b0: k = kq, kq + 1
                    o[2 * h + 1] = at(k0 + kq + 8, n) | (at(k0 + kq + 9, n) << 16);       // b1: k + 8
                }
            }
}

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t &a0, uint32_t &a1, uint32_t &a2, uint32_t &a3)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(a0), "=r"(a1), "=r"(a2), "=r"(a3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// the weights of k-pairs kp0 .. kp0 + NKP of one n-tile (w = the tile's first uint4 for lane 0), issued together
template <int NKP>
__device__ __forceinline__ void tail_load(uint4 (&bw)[NKP], const uint4 *w, int kp0, int lane)
{
#pragma unroll
    for (int i = 0; i < NKP; ++i) bw[i] = __ldg(w + (size_t)(kp0 + i) * 32 + lane);
}
// acc[mt] += A[32 rows, those k-pairs] . W[n-tile]; A in shared memory (row stride LD bf16)
template <int LD, int NKP>
__device__ __forceinline__ void tail_mma(const __nv_bfloat16 *A, const uint4 (&bw)[NKP], int kp0, float (&acc)[2][4], int lane)
{
    const uint32_t a_base = (uint32_t)__cvta_generic_to_shared(A) + (uint32_t)(((lane & 15) * LD + (lane >> 4) * 8) * 2);
#pragma unroll
    for (int i = 0; i < NKP; ++i) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int k0 = ((kp0 + i) * 2 + h) * 16;
            const uint32_t b0 = h ? bw[i].z : bw[i].x, b1 = h ? bw[i].w : bw[i].y;
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                uint32_t a0, a1, a2, a3;
                ldmatrix_x4(a_base + (uint32_t)((mt * 16 * LD + k0) * 2), a0, a1, a2, a3);
                mma_bf16(acc[mt], a0, a1, a2, a3, b0, b1);
            }
        }
    }
}

struct TailShared {
    __nv_bfloat16 A1[kTailMaxRows * kLd1], A2[kTailMaxRows * kLd2], A3[kTailMaxRows * kLd3];
    float red[6][2][32][4];                              // partial accumulators of the warps that do not finish a tile
    __nv_bfloat16 stage[kTailMaxRows * 32];              // this CTA's slice of a layer's output before it is pushed to the peers
    float logits[kTailMaxRows][33];
};

// bias + optional ReLU of a finished 32 x 8 tile -> bf16 into stage[row][col0 + ...] (row stride `sld`)
__device__ __forceinline__ void tail_store_stage(__nv_bfloat16 *stage, int sld, int col0, const float (&acc)[2][4], const float *bias, int lane)
{
    const int g = lane >> 2, q = (lane & 3) * 2;
    const float t0 = __ldg(bias + q), t1 = __ldg(bias + q + 1);
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int row = mt * 16 + hh * 8 + g;
            const float v0 = fmaxf(acc[mt][2 * hh] + t0, 0.f), v1 = fmaxf(acc[mt][2 * hh + 1] + t1, 0.f);
            *reinterpret_cast<__nv_bfloat162 *>(stage + row * sld + col0 + q) = __floats2bfloat162_rn(v0, v1);
        }
}

__global__ void __cluster_dims__(kTailCluster, 1, 1) __launch_bounds__(kTailThreads, 1)
tail_fused_kernel(TailFusedParams p)
{
    extern __shared__ __align__(16) unsigned char tail_smem[];
    TailShared &S = *reinterpret_cast<TailShared *>(tail_smem);
    cg::cluster_group cluster = cg::this_cluster();
    const int c = (int)cluster.block_rank();
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;

    // programmatic dependent launch: the successor may start its own prologue now; this kernel's weights do not depend on the
    // predecessor, its activations do
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int nt1 = w & 3, kh1 = w >> 2;            // layer 1: this CTA's 32 columns = 4 n-tiles x 2 K halves over the 8 warps
    const int nt2 = w & 1, kq2 = w >> 1;            // layer 2: 16 columns = 2 n-tiles x 4 K quarters
    const int NT3 = (p.n3 + 7) / 8;
    const int t3 = NT3 <= 4 ? w : w * kTailCluster + c;      // layer 3: first n-tile of this warp (narrow: CTA 0 only)
    uint4 bw1[16], bw2[4], bw3[8];
    tail_load<16>(bw1, p.w1 + (size_t)(4 * c + nt1) * (1024 / 32) * 32, kh1 * 16, lane);
    tail_load<4>(bw2, p.w2 + (size_t)(2 * c + nt2) * (512 / 32) * 32, kq2 * 4, lane);
    if (t3 < NT3) tail_load<8>(bw3, p.w3 + (size_t)t3 * (256 / 32) * 32, 0, lane);
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // ---- A1 = bf16(pooled), rows >= b zero
#pragma unroll 8
    for (int idx = tid; idx < kTailMaxRows * 256; idx += kTailThreads) {
        const int row = idx >> 8, c4 = idx & 255;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < p.b) v = __ldg(reinterpret_cast<const float4 *>(p.A0 + (size_t)row * 1024) + c4);
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 u;
        u.x = *reinterpret_cast<uint32_t *>(&lo); u.y = *reinterpret_cast<uint32_t *>(&hi);
        *reinterpret_cast<uint2 *>(S.A1 + row * kLd1 + c4 * 4) = u;
    }
    cluster.sync();               // A1 complete, and every CTA of the cluster is running before anybody writes into a peer

    // ---- layer 1: 1024 -> 512
    {
        const int nt = nt1, kh = kh1;
        float acc[2][4] = {};
        tail_mma<kLd1, 16>(S.A1, bw1, kh * 16, acc, lane);
        if (kh == 1) {
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int i = 0; i < 4; ++i) S.red[nt][mt][lane][i] = acc[mt][i];
        }
        __syncthreads();
        if (kh == 0) {
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[mt][i] += S.red[nt][mt][lane][i];
            tail_store_stage(S.stage, 32, nt * 8, acc, p.t1 + 32 * c + nt * 8, lane);
        }
        __syncthreads();
        // push the 32 x 32 slice into every CTA's A2 (columns 32 c ..): 128 uint4 per peer
        for (int idx = tid; idx < kTailCluster * 128; idx += kTailThreads) {
            const int peer = idx >> 7, q = idx & 127, row = q >> 2, part = q & 3;
            const uint4 v = *reinterpret_cast<const uint4 *>(S.stage + row * 32 + part * 8);
            *reinterpret_cast<uint4 *>(cluster.map_shared_rank(S.A2, peer) + row * kLd2 + 32 * c + part * 8) = v;
        }
        cluster.sync();
    }
    // ---- layer 2: 512 -> 256
    {
        const int nt = nt2, kq = kq2;
        float acc[2][4] = {};
        tail_mma<kLd2, 4>(S.A2, bw2, kq * 4, acc, lane);
        if (kq > 0) {
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int i = 0; i < 4; ++i) S.red[(kq - 1) * 2 + nt][mt][lane][i] = acc[mt][i];
        }
        __syncthreads();
        if (kq == 0) {
#pragma unroll
            for (int j = 0; j < 3; ++j)
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                    for (int i = 0; i < 4; ++i) acc[mt][i] += S.red[j * 2 + nt][mt][lane][i];
            tail_store_stage(S.stage, 16, nt * 8, acc, p.t2 + 16 * c + nt * 8, lane);
        }
        __syncthreads();
        for (int idx = tid; idx < kTailCluster * 64; idx += kTailThreads) {
            const int peer = idx >> 6, q = idx & 63, row = q >> 1, part = q & 1;
            const uint4 v = *reinterpret_cast<const uint4 *>(S.stage + row * 16 + part * 8);
            *reinterpret_cast<uint4 *>(cluster.map_shared_rank(S.A3, peer) + row * kLd3 + 16 * c + part * 8) = v;
        }
        cluster.sync();           // the last remote access to any peer's shared memory
    }
    // ---- layer 3: 256 -> n3 (no ReLU)
    const int g = lane >> 2, q = (lane & 3) * 2;
    if (NT3 <= 4) {
        // narrow output (R, logits): CTA 0 alone, one warp per n-tile, then the row softmax if asked for
        if (c != 0) return;
        if (w < NT3) {
            float acc[2][4] = {};
            tail_mma<kLd3, 8>(S.A3, bw3, 0, acc, lane);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt)
#pragma unroll
                for (int hh = 0; hh < 2; ++hh)
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int col = w * 8 + q + e;
                        if (col < p.n3) S.logits[mt * 16 + hh * 8 + g][col] = acc[mt][2 * hh + e] + __ldg(p.t3 + col);
                    }
        }
        __syncthreads();
        if (tid < p.b) {
            const float *l = S.logits[tid];
            float mx = -INFINITY;
            for (int j = 0; j < p.n3; ++j) mx = fmaxf(mx, l[j]);
            float sum = 0.f;
            for (int j = 0; j < p.n3; ++j) sum += expf(l[j] - mx);
            const float inv = 1.f / sum;
            for (int j = 0; j < p.n3; ++j) {
                if (p.out32) p.out32[(size_t)tid * p.n3 + j] = l[j];
                if (p.out16) p.out16[(size_t)tid * p.n3 + j] = __float2bfloat16_rn(l[j]);
                if (p.out_softmax) p.out_softmax[(size_t)tid * p.n3 + j] = expf(l[j] - mx) * inv;
            }
        }
        return;
    }
    // wide output (the folded per-cloud operand image, 4096 columns): n-tiles dealt over CTAs and warps, results straight to HBM;
    // the next tile's weights are in flight while this one is multiplied
    for (int t = t3; t < NT3; t += 8 * kTailCluster) {
        uint4 nx[8];
        const int tn = t + 8 * kTailCluster;
        if (tn < NT3) tail_load<8>(nx, p.w3 + (size_t)tn * (256 / 32) * 32, 0, lane);
        float acc[2][4] = {};
        tail_mma<kLd3, 8>(S.A3, bw3, 0, acc, lane);
        const int col = t * 8 + q;
        const float t0 = col < p.n3 ? __ldg(p.t3 + col) : 0.f, t1 = col + 1 < p.n3 ? __ldg(p.t3 + col + 1) : 0.f;
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int row = mt * 16 + hh * 8 + g;
                if (row >= p.b) continue;
                const float v0 = acc[mt][2 * hh] + t0, v1 = acc[mt][2 * hh + 1] + t1;
                if (col + 1 < p.n3) {
                    if (p.out32) *reinterpret_cast<float2 *>(p.out32 + (size_t)row * p.n3 + col) = make_float2(v0, v1);
                    if (p.out16) *reinterpret_cast<__nv_bfloat162 *>(p.out16 + (size_t)row * p.n3 + col) = __floats2bfloat162_rn(v0, v1);
                } else if (col < p.n3) {
                    if (p.out32) p.out32[(size_t)row * p.n3 + col] = v0;
                    if (p.out16) p.out16[(size_t)row * p.n3 + col] = __float2bfloat16_rn(v0);
                }
            }
        if (tn < NT3) {
#pragma unroll
            for (int i = 0; i < 8; ++i) bw3[i] = nx[i];
        }
    }
}

int tail_fused(const TailFusedParams &p, cudaStream_t st)
{
    PCP_REQUIRE(p.A0 && p.w1 && p.w2 && p.w3 && p.t1 && p.t2 && p.t3 && p.b >= 1 && p.b <= kTailMaxRows && p.n3 >= 1, PCP_ERR_INVALID,
                "tail_fused: bad arguments (b=%d n3=%d)", p.b, p.n3);
    PCP_REQUIRE(!p.out_softmax || p.n3 <= 32, PCP_ERR_INVALID, "tail_fused: softmax needs n3 <= 32");
    PCP_REQUIRE(p.n3 <= 32 || p.n3 % 2 == 0, PCP_ERR_INVALID, "tail_fused: a wide third layer needs an even column count");
    static bool attr_set[64] = {};
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    PCP_REQUIRE(dev >= 0 && dev < 64, PCP_ERR_UNSUPPORTED, "tail_fused: device ordinal %d", dev);
    if (!attr_set[dev]) {
        PCP_CUDA_OK(cudaFuncSetAttribute(tail_fused_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        PCP_CUDA_OK(cudaFuncSetAttribute(tail_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TailShared)));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kTailCluster);
    cfg.blockDim = dim3(kTailThreads);
    cfg.dynamicSmemBytes = sizeof(TailShared);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;      // starts while the producer of A0 drains; see griddepcontrol.wait
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, tail_fused_kernel, p));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

}  // namespace pcp
