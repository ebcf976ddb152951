// tc_gemm_tma.cu — the large-M form of tc_gemm (per-point layers: segmentation head, training forward / input
// gradients): a PERSISTENT kernel, one CTA per SM, that walks 128 x BNt output tiles.
//   * activation tiles (bf16, row-major in HBM) arrive by TMA tensor copies (cp.async.bulk.tensor.2d, SWIZZLE_128B box of
//     128 rows x 64 K) straight into the UMMA operand layout: no registers, four 16 KB tiles in flight per SM;
//   * weight tiles come pre-packed (pack_gemm_b_image) by cp.async.bulk on the same stage barrier;
//   * one thread issues tcgen05.mma into one of two 128-column TMEM accumulators;
//   * kPGroups epilogue warpgroups take tiles round-robin (accumulator i % kPGroups): TMEM -> registers -> bias / ReLU -> bf16 or fp32 -> a
//     warp-private shared tile -> 128-byte-per-row coalesced global stores, overlapping the next tile's loads and MMAs.
// The per-CTA launch / TMEM-alloc / barrier-init cost of the one-tile-per-CTA kernel (4-5 us per tile, measured with its
// loads and stores compiled out) is paid once per SM instead of once per tile.
#include <cuda.h>

#include "tc_gemm.cuh"
#include "umma.cuh"

namespace pcp {

using namespace umma;

constexpr int kPGroups = 2;                // epilogue warpgroups = TMEM accumulators in flight (the epilogue is latency-bound: more warps)
constexpr int kPThreads = (4 * kPGroups + 2) * 32;   // epilogue groups, then the loader warp, then the MMA issuer warp
constexpr int kPLoaderWarp = 4 * kPGroups, kPMmaWarp = 4 * kPGroups + 1;
constexpr uint32_t kPTmemCols = 256;       // kPGroups x 128 rounded up to a power of two
constexpr int kPMaxStages = 8;
constexpr uint32_t kPStageTile = 32 * 144; // per-warp staging tile: 32 rows x (128 B + 16 B pad)
constexpr uint32_t kPOperandBytes = 180224; // 176 KB of operand space: streaming = 5 stages x (16 KB A + 16 KB B); resident = B image + A ring
constexpr uint32_t kPOffStage = kPOperandBytes;
constexpr uint32_t kPOffBias = kPOffStage + 4 * kPGroups * kPStageTile;
constexpr uint32_t kPSmem = kPOffBias + kPGroups * 1024 + 1024;

// Shared-memory plan.  resident: the CTA's weight column block (KT tiles) sits at offset 0 for the whole kernel and the
// activation ring follows; streaming: each stage holds an activation tile and the weight tile of the same k-block.
struct PPlan {
    int resident, stages;
    uint32_t a_off0, a_stride, b_off0, b_stride;       // A tile of stage s: a_off0 + s * a_stride; B tile: b_off0 + (resident ? kt : s) * b_stride
};

__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *tm, int c0, int c1, uint64_t *bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 :: "r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}

// (v0 + b.x, v1 + b.y) as one packed fp32x2 add, then ONE conversion to a bf16 pair with the ReLU folded into it:
// two instructions per two outputs instead of five (FADD, FADD, FMNMX, FMNMX, F2FP) — the epilogue is issue-bound
template <bool kRelu>
__device__ __forceinline__ uint32_t bias_act_bf16x2(uint32_t v0, uint32_t v1, float2 b)
{
    uint64_t a, bb, r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "r"(v0), "r"(v1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b.x), "f"(b.y));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(bb));
    float lo, hi;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(r));
    uint32_t d;
    if (kRelu) asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    else asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}

template <bool kRelu>
__device__ __forceinline__ void stage_bf16_chunk(const uint32_t (&v)[32], const float *bias, uint8_t *dst_row)
{
#pragma unroll
    for (int q = 0; q < 32; q += 8) {
        const float4 b0 = *reinterpret_cast<const float4 *>(bias + q), b1 = *reinterpret_cast<const float4 *>(bias + q + 4);
        *reinterpret_cast<uint4 *>(dst_row + q * 2) =
            make_uint4(bias_act_bf16x2<kRelu>(v[q], v[q + 1], make_float2(b0.x, b0.y)), bias_act_bf16x2<kRelu>(v[q + 2], v[q + 3], make_float2(b0.z, b0.w)),
                       bias_act_bf16x2<kRelu>(v[q + 4], v[q + 5], make_float2(b1.x, b1.y)), bias_act_bf16x2<kRelu>(v[q + 6], v[q + 7], make_float2(b1.z, b1.w)));
    }
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory"); }

__global__ void __launch_bounds__(kPThreads, 1)
tc_gemm_persist_kernel(const __grid_constant__ CUtensorMap tmA, const TcGemmParams p, const PPlan pl, int bnt, int mtiles, int ntiles, int total_tiles)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full[kPMaxStages], empty[kPMaxStages], acc_full[kPGroups], acc_empty[kPGroups], b_full;
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int KT = p.K >> 6;
    const uint32_t b_tile_bytes = (uint32_t)bnt * 128u;

    if (tid == 0) {
        for (int i = 0; i < kPMaxStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < kPGroups; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
        mbar_init(&b_full, 1);
        fence_barrier_init();
    }
    if (warp == kPMmaWarp) tmem_alloc(&tmem_base, kPTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // programmatic dependent launch: see tc_gemm.cu
    // tile walk.  streaming: t = blockIdx.x, += gridDim.x over (gm, nt) with nt fastest.  resident: this CTA keeps column block
    // nt_fix = blockIdx.x % ntiles and walks gm = blockIdx.x / ntiles, += gridDim.x / ntiles (the host makes gridDim.x a multiple
    // of ntiles); neighbouring CTAs read the same activation tile at the same time, so HBM sees it once.
    const int t_first = pl.resident ? (blockIdx.x / ntiles) * ntiles + (blockIdx.x % ntiles) : (int)blockIdx.x;
    const int t_step = pl.resident ? ((int)gridDim.x / ntiles) * ntiles : (int)gridDim.x;
    const uint32_t nstage = (uint32_t)pl.stages;

    if (warp == kPLoaderWarp) {
        // ---------------- loader: one thread, TMA for activations + bulk copy for the weight tile ----------------
        if (lane == 0) {
            uint32_t it = 0;
            if (pl.resident) {
                const uint8_t *img = p.Bimg + (size_t)(blockIdx.x % ntiles) * KT * b_tile_bytes;
                mbar_arrive_expect_tx(&b_full, (uint32_t)KT * b_tile_bytes);
                for (int kt = 0; kt < KT; ++kt) bulk_g2s(base + pl.b_off0 + kt * pl.b_stride, img + (size_t)kt * b_tile_bytes, b_tile_bytes, &b_full);
            }
            const bool lda = !(p.dbg & 2);
            asm volatile("griddepcontrol.wait;" ::: "memory");       // activations are the predecessor's output; the weights above are not
            for (int t = t_first; t < total_tiles; t += t_step) {
                const int nt = t % ntiles, gm = t / ntiles;
                const int g = gm / mtiles, mt = gm % mtiles;
                const int arow = g * p.M + mt * 128;
                const uint8_t *img = p.Bimg + (size_t)g * p.b_group_stride + (size_t)nt * KT * b_tile_bytes;
                for (int kt = 0; kt < KT; ++kt, ++it) {
                    const uint32_t s = it % nstage;
                    if (it >= nstage) mbar_wait(&empty[s], ((it / nstage) - 1) & 1, 50);
                    mbar_arrive_expect_tx(&full[s], (pl.resident ? 0u : b_tile_bytes) + (lda ? 16384u : 0u));
                    if (lda) tma_load_2d(base + pl.a_off0 + s * pl.a_stride, &tmA, kt * 64, arow, &full[s]);
                    if (!pl.resident) bulk_g2s(base + pl.b_off0 + s * pl.b_stride, img + (size_t)kt * b_tile_bytes, b_tile_bytes, &full[s]);
                }
            }
        }
    } else if (warp == kPMmaWarp) {
        // ---------------- MMA issuer ----------------
        const uint32_t idesc = make_idesc_bf16(128, (uint32_t)bnt);
        const uint32_t lo_base = (smem_u32(base) & 0x3FFFFu) >> 4;
        uint32_t it = 0, i = 0;
        if (pl.resident) mbar_wait(&b_full, 0, 54);
        for (int t = t_first; t < total_tiles; t += t_step, ++i) {
            const uint32_t buf = i % kPGroups;
            if (i >= (uint32_t)kPGroups) mbar_wait(&acc_empty[buf], ((i / kPGroups) - 1) & 1, 51);
            tc_fence_after();
            const uint32_t d = tmem + buf * 128u;
            for (int kt = 0; kt < KT; ++kt, ++it) {
                const uint32_t s = it % nstage;
                mbar_wait(&full[s], (it / nstage) & 1, 52);
                tc_fence_after();
                const uint32_t a_lo = lo_base + ((pl.a_off0 + s * pl.a_stride) >> 4);
                const uint32_t b_lo = lo_base + ((pl.b_off0 + (pl.resident ? (uint32_t)kt : s) * pl.b_stride) >> 4);
                if (elect_one_sync()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) mma_bf16_ss(d, sdesc_from_lo(a_lo + k * 2), sdesc_from_lo(b_lo + k * 2), idesc, (kt | k) > 0);
                    mma_commit(&empty[s]);
                    if (kt == KT - 1) mma_commit(&acc_full[buf]);
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------- epilogue groups: group e (warps 4e..4e+3) takes the tiles with local index i % 2 == e ----------------
        // (all eight warps on every tile was measured slower: one group's arithmetic and stores overlap the other's TMEM reads)
        asm volatile("griddepcontrol.wait;" ::: "memory");           // bias may be the predecessor's output; C may still be read by it
        const int e = warp >> 2, wq = warp & 3, gt = tid & 127;
        uint8_t *stage = base + kPOffStage + warp * kPStageTile;
        float *bias_buf = reinterpret_cast<float *>(base + kPOffBias + e * 1024);     // two 128-float tiles per group
        const float *bias_s = bias_buf;
        int bias_key = -1, bias_sel = 1;
        uint32_t i = 0;
        for (int t = t_first; t < total_tiles; t += t_step, ++i) {
            if ((int)(i % kPGroups) != e) continue;
            const int nt = t % ntiles, gm = t / ntiles;
            const int g = gm / mtiles, mt = gm % mtiles;
            const int row0 = mt * 128;
            const int rows_valid = min(128, p.M - row0);
            const int ncol0 = nt * bnt;
            // bias tile of this (group, column block): reloaded only when it changes, into the OTHER buffer — every thread of the
            // group has left the previous tile when it reaches the barrier, so the buffer of two changes ago is free
            const int key = (p.bias_group_stride ? g : 0) * ntiles + nt;
            if (key != bias_key) {
                bias_key = key;
                bias_sel ^= 1;
                const int n = ncol0 + gt;
                bias_buf[bias_sel * 128 + gt] = (p.bias && n < p.N) ? __ldg(p.bias + (size_t)g * p.bias_group_stride + n) : 0.f;
                named_bar_sync(1 + e, 128);
                bias_s = bias_buf + bias_sel * 128;
            }
            // bf16-only output with whole, aligned 32-column chunks: the short path below
            const bool fast16 = p.C16 && !p.C32 && !p.Csoftmax && ncol0 + bnt <= p.N && (bnt & 31) == 0 && (p.ldc & 7) == 0 &&
                                ((reinterpret_cast<uintptr_t>(p.C16) & 15u) == 0) && (p.c_group_stride & 7) == 0;
            // fp32-only output with whole, aligned chunks (the training step's Z / dY): its own short path as well — through the
            // generic code below the kernel spent 45 % of its stall samples waiting for INSTRUCTIONS (ncu: stall_no_inst; the
            // four inlined copies of every output variant overflow the instruction cache), 36 us for 2048 tiles with all memory
            // traffic compiled out against 15 us for the bf16 path
            const bool fast32 = p.C32 && !p.C16 && !p.Csoftmax && ncol0 + bnt <= p.N && (bnt & 31) == 0 && (p.ldc & 3) == 0 &&
                                ((reinterpret_cast<uintptr_t>(p.C32) & 15u) == 0) && (p.c_group_stride & 3) == 0;
            const int rv = (p.dbg & 1) ? 0 : rows_valid - wq * 32;                      // rows of this warp's quarter that exist
            mbar_wait(&acc_full[e], (i / kPGroups) & 1, 53);
            tc_fence_after();
            const uint32_t trow = tmem + e * 128u + ((uint32_t)(wq * 32) << 16);
            const size_t c_row0 = (size_t)g * p.c_group_stride + (size_t)(row0 + wq * 32) * p.ldc;
            // the accumulator row is read 64 columns at a time (two loads in flight, one wait — a thread's budget is 168 registers
            // because the 10-warp CTA is allocated as 12), and handed back to the MMA issuer right after the last read
#pragma unroll
            for (int half = 0; half < 2; ++half) {
            if (half * 64 >= bnt) break;
            uint32_t vv[2][32];
#pragma unroll
            for (int c = 0; c < 2; ++c) if ((2 * half + c) * 32 < bnt) tmem_ld32(trow + (2 * half + c) * 32, vv[c]);
            tmem_ld_wait();
            if ((half + 1) * 64 >= bnt) {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[e]);
            }
#pragma unroll
            for (int c = 0; c < 2; ++c) {
                const int c0 = (2 * half + c) * 32;
                if (c0 >= bnt) break;
                const uint32_t (&v)[32] = vv[c];
                if (fast16) {
                    if (p.relu) stage_bf16_chunk<true>(v, bias_s + c0, stage + lane * 144);
                    else stage_bf16_chunk<false>(v, bias_s + c0, stage + lane * 144);
                    __syncwarp();
                    __nv_bfloat16 *dst = p.C16 + c_row0 + (size_t)(lane >> 2) * p.ldc + ncol0 + c0 + (lane & 3) * 8;
                    const uint8_t *src = stage + (lane >> 2) * 144 + (lane & 3) * 16;
#pragma unroll
                    for (int r8 = 0; r8 < 4; ++r8)
                        if (r8 * 8 + (lane >> 2) < rv)
                            *reinterpret_cast<uint4 *>(dst + (size_t)(r8 * 8) * p.ldc) = *reinterpret_cast<const uint4 *>(src + r8 * 8 * 144);
                    __syncwarp();
                    continue;
                }
                if (fast32) {
                    const float *bs = bias_s + c0;
#pragma unroll
                    for (int q = 0; q < 32; q += 4) {
                        const float4 bq = *reinterpret_cast<const float4 *>(bs + q);
                        float4 o4 = make_float4(__uint_as_float(v[q]) + bq.x, __uint_as_float(v[q + 1]) + bq.y, __uint_as_float(v[q + 2]) + bq.z,
                                                __uint_as_float(v[q + 3]) + bq.w);
                        if (p.relu) { o4.x = fmaxf(o4.x, 0.f); o4.y = fmaxf(o4.y, 0.f); o4.z = fmaxf(o4.z, 0.f); o4.w = fmaxf(o4.w, 0.f); }
                        *reinterpret_cast<float4 *>(stage + lane * 144 + q * 4) = o4;
                    }
                    __syncwarp();
                    // one store instruction = four rows x 128 contiguous bytes
                    float *dst = p.C32 + c_row0 + (size_t)(lane >> 3) * p.ldc + ncol0 + c0 + (lane & 7) * 4;
                    const uint8_t *src = stage + (lane >> 3) * 144 + (lane & 7) * 16;
#pragma unroll
                    for (int r4 = 0; r4 < 8; ++r4)
                        if (r4 * 4 + (lane >> 3) < rv) {
                            float4 q4 = *reinterpret_cast<const float4 *>(src + r4 * 4 * 144);
                            float4 *d4 = reinterpret_cast<float4 *>(dst + (size_t)(r4 * 4) * p.ldc);
                            if (p.accumulate) { const float4 old = *d4; q4.x += old.x; q4.y += old.y; q4.z += old.z; q4.w += old.w; }
                            *d4 = q4;
                        }
                    __syncwarp();
                    continue;
                }
                float o[32];
#pragma unroll
                for (int q = 0; q < 32; q += 4) {
                    const float4 bq = *reinterpret_cast<const float4 *>(bias_s + c0 + q);
                    o[q] = __uint_as_float(v[q]) + bq.x; o[q + 1] = __uint_as_float(v[q + 1]) + bq.y;
                    o[q + 2] = __uint_as_float(v[q + 2]) + bq.z; o[q + 3] = __uint_as_float(v[q + 3]) + bq.w;
                }
                if (p.relu) {
#pragma unroll
                    for (int q = 0; q < 32; ++q) o[q] = fmaxf(o[q], 0.f);
                }
                const int nbase = ncol0 + c0;
                if (p.Csoftmax) {                          // N <= 32: the whole row is in this chunk; thread = row
                    float mx = -3.0e38f;
#pragma unroll
                    for (int q = 0; q < 32; ++q) if (q < p.N) mx = fmaxf(mx, o[q]);
                    float sum = 0.f;
#pragma unroll
                    for (int q = 0; q < 32; ++q) { o[q] = (q < p.N) ? __expf(o[q] - mx) : 0.f; sum += o[q]; }
                    const float inv = 1.0f / sum;
                    if (wq * 32 + lane < rows_valid && !(p.dbg & 1)) {
                        float *dst = p.Csoftmax + c_row0 + (size_t)lane * p.ldc;
#pragma unroll
                        for (int q = 0; q < 32; ++q) if (q < p.N) dst[q] = o[q] * inv;
                    }
                    continue;
                }
                if (p.C16) {
                    // 32 columns = 64 B per row: stage, then each store instruction writes 8 rows x 64 contiguous bytes
#pragma unroll
                    for (int q = 0; q < 32; q += 8)
                        *reinterpret_cast<uint4 *>(stage + lane * 144 + q * 2) = make_uint4(pack_bf16x2(o[q], o[q + 1]), pack_bf16x2(o[q + 2], o[q + 3]),
                                                                                           pack_bf16x2(o[q + 4], o[q + 5]), pack_bf16x2(o[q + 6], o[q + 7]));
                    __syncwarp();
                    const bool fast = nbase + 32 <= p.N && (p.ldc & 7) == 0 && ((reinterpret_cast<uintptr_t>(p.C16) & 15u) == 0) && (p.c_group_stride & 7) == 0;
#pragma unroll
                    for (int r8 = 0; r8 < 4; ++r8) {
                        const int rr = r8 * 8 + (lane >> 2), cc = (lane & 3) * 8;
                        if (wq * 32 + rr < rows_valid && !(p.dbg & 1)) {
                            const uint4 q4 = *reinterpret_cast<const uint4 *>(stage + rr * 144 + cc * 2);
                            __nv_bfloat16 *dst = p.C16 + c_row0 + (size_t)rr * p.ldc + nbase + cc;
                            if (fast) *reinterpret_cast<uint4 *>(dst) = q4;
                            else {
                                const __nv_bfloat16 *h = reinterpret_cast<const __nv_bfloat16 *>(&q4);
                                for (int z = 0; z < 8; ++z) if (nbase + cc + z < p.N) dst[z] = h[z];
                            }
                        }
                    }
                    __syncwarp();
                }
                if (p.C32) {
#pragma unroll
                    for (int q = 0; q < 32; q += 4) *reinterpret_cast<float4 *>(stage + lane * 144 + q * 4) = make_float4(o[q], o[q + 1], o[q + 2], o[q + 3]);
                    __syncwarp();
                    const bool fast = nbase + 32 <= p.N && (p.ldc & 3) == 0 && ((reinterpret_cast<uintptr_t>(p.C32) & 15u) == 0) && (p.c_group_stride & 3) == 0;
#pragma unroll
                    for (int r4 = 0; r4 < 8; ++r4) {
                        const int rr = r4 * 4 + (lane >> 3), cc = (lane & 7) * 4;
                        if (wq * 32 + rr < rows_valid && !(p.dbg & 1)) {
                            float4 q4 = *reinterpret_cast<const float4 *>(stage + rr * 144 + cc * 4);
                            float *dst = p.C32 + c_row0 + (size_t)rr * p.ldc + nbase + cc;
                            if (fast) {
                                if (p.accumulate) {
                                    const float4 old = *reinterpret_cast<const float4 *>(dst);
                                    q4.x += old.x; q4.y += old.y; q4.z += old.z; q4.w += old.w;
                                }
                                *reinterpret_cast<float4 *>(dst) = q4;
                            } else {
                                const float *h = reinterpret_cast<const float *>(&q4);
                                for (int z = 0; z < 4; ++z) if (nbase + cc + z < p.N) dst[z] = p.accumulate ? dst[z] + h[z] : h[z];
                            }
                        }
                    }
                    __syncwarp();
                }
            }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kPMmaWarp) tmem_dealloc(tmem, kPTmemCols);
}

// ---------------------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn()
{
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
        else
            cudaGetLastError();
    }
    return fn;
}

// Returns true when the persistent TMA kernel took the call (*rc holds its status); false = caller uses the per-tile kernel.
bool tc_gemm_persist_try(const TcGemmParams &p, cudaStream_t st, int *rc)
{
    static const bool off = getenv("PCP_TC_GEMM_NO_PERSIST") != nullptr;
    if (off || p.a_f32 || p.M <= 0 || p.groups <= 0) return false;
    if (p.groups > 1 && p.a_group_stride != (int64_t)p.M * p.lda) return false;
    if ((p.lda & 7) || (reinterpret_cast<uintptr_t>(p.A) & 15u)) return false;
    const int bnt = tc_gemm_bn_tile(p.N);
    const int ntiles = (p.N + bnt - 1) / bnt, mtiles = (p.M + 127) / 128;
    const int64_t total = (int64_t)ntiles * mtiles * p.groups;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (total < 2 * (int64_t)sms || total > 0x7fffffff) return false;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    CUtensorMap tm;
    const cuuint64_t dims[2] = {(cuuint64_t)p.K, (cuuint64_t)p.M * (cuuint64_t)p.groups};
    const cuuint64_t strides[1] = {(cuuint64_t)p.lda * 2};
    const cuuint32_t box[2] = {64, 128}, estr[2] = {1, 1};
    if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(p.A), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return false;
    static bool attr = false;
    if (!attr) {
        if (cudaFuncSetAttribute(tc_gemm_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPSmem) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        attr = true;
    }
    // shared-memory plan: keep the CTA's weight column block resident when it fits beside a >= 6-deep activation ring (a shallower ring starves the MMA queue: measured)
    const int KT = p.K / 64;
    const uint32_t b_tile = (uint32_t)bnt * 128u, b_tile_al = (b_tile + 1023u) & ~1023u;
    PPlan pl;
    int grid = (int)std::min<int64_t>(total, sms);
    const uint32_t rb = (uint32_t)KT * b_tile_al;
    if (p.b_group_stride == 0 && rb + 6u * 16384u <= kPOperandBytes && grid >= ntiles) {
        pl.resident = 1;
        pl.b_off0 = 0; pl.b_stride = b_tile_al;
        pl.a_off0 = rb; pl.a_stride = 16384;
        pl.stages = (int)std::min<uint32_t>(kPMaxStages, (kPOperandBytes - rb) / 16384u);
        grid = (grid / ntiles) * ntiles;
    } else {
        pl.resident = 0;
        pl.a_off0 = 0; pl.a_stride = 32768; pl.b_off0 = 16384; pl.b_stride = 32768;
        pl.stages = (int)(kPOperandBytes / 32768u);
    }
    static const bool no_pdl = getenv("PCP_TC_GEMM_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kPThreads); cfg.dynamicSmemBytes = kPSmem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = no_pdl ? 0 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, tc_gemm_persist_kernel, tm, p, pl, bnt, mtiles, ntiles, (int)total);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("tc_gemm_persist: %s", cudaGetErrorString(e)); *rc = PCP_ERR_CUDA; }
    else *rc = PCP_OK;
    return true;
}

}  // namespace pcp
