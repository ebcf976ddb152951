// chain_tc.cu — the per-point MLP chains of PointNet (pointnet/PointNet.py:215-248, TNet.call :423-429)
// as ONE fused tcgen05 kernel per chain: raw points -> normalise -> 3->64 (CUDA-core FMA) ->
// 64->64/128 mid layers (tcgen05.mma, accumulators in TMEM, BN-folded bias + ReLU epilogue writing the
// next layer's bf16 operand straight into swizzled shared memory) -> 128->1024 final layer evaluated
// TRANSPOSED (D[channel, point] = W3^T . H2^T) so the global max-pool over points is a pure in-thread
// FMNMX over TMEM columns -> per-cloud atomicMax.  No per-point activation ever reaches HBM: the chain
// reads 12 B/point and writes 4 KB/cloud.
//
// Warp roles (512 threads, one CTA per SM, CTA pairs (cta_group::2), persistent over a contiguous range of pair-steps of
// 2 x 128 points, TWO pair-steps in flight — even steps belong to front group 0, odd steps to front group 1):
//   warps 0-3   front group 0: normalise + L0 + mid-layer epilogues (TMEM -> regs -> bf16 -> smem, IN PLACE into the operand
//               buffer A0 its own MMA has finished reading; the last layer's epilogue writes the shared H2 halves)
//   warps 12-15 front group 1: the same on buffer A1 / the second mid accumulator (same TMEM lane quarters as warps 0-3)
//   warps 4-7   final epilogue: TMEM (two 64-column passes) -> running per-channel max -> atomicMax
//   warp  8/11  mid-layer MMA issuer of group 0 / 1 (one elected thread each)
//   warp  10    final-layer MMA issuer (two granules in flight)
//   warp  9     loader (one thread): weight images via cp.async.bulk (TMA engine) + TMEM allocation
// TMEM: 2 x 128 columns mid accumulators + 2 x 128 columns final accumulators.  Barriers: a_full[g] / mid_full[g] per group,
// h2_full[half], h2_empty[step parity][half] (a parity wait is only alias-free when the waiter is at most one phase behind).
#include <cstring>
#include <vector>

#include "gemm_simt.cuh"
#include "model.cuh"
#include "umma.cuh"

namespace pcp {

using namespace umma;

constexpr int kTP = 128;            // points per tile
constexpr int kChainThreads = 512;   // 16 warps: front group 0 (0-3), final epilogue (4-7), mid issuer 0 (8), loader (9), final issuer (10), mid issuer 1 (11), front group 1 (12-15)
constexpr uint32_t kW3Tile = 16384;  // one W3 operand tile: 128 channels x 64 K, bf16

struct ChainParams {
    const float *x; int64_t n; int b;
    const float *centroid, *scale;        // (b,3), (b)
    const float *R;                       // (b,9) or nullptr: X = pcn @ R before L0 (PointNet.py:207)
    const float *l0;                      // [64][4]: w0, w1, w2, t (BN folded)
    int nmid;                             // mid layers (all K = 64)
    int mid_n[3];                         // output widths; the last one is 128
    int mid_slot[3];                      // smem weight slot
    const uint8_t *mid_w[3];              // bf16 SW128 images (rows = mid_n, K = 64)
    int64_t mid_cloud_stride[3];          // bytes between clouds for a per-cloud layer, 0 = shared
    const float *mid_t[3];
    const uint8_t *w3;                    // 16 tiles (j = channel block, h = K half) of 16 KB
    const float *t3;                      // 1024
    float *gmax;                          // (b,1024), zero-initialised
    int tiles_per_cloud;
    int64_t total_tiles;
    unsigned long long *trace;            // development only (PCP_CHAIN_TRACE): per-role event timestamps of cluster 0
    int trace_tid;                        // the front-group thread that records role 0 (PCP_CHAIN_TRACE_TID, default 0)
};

// ---------------------------------------------------------------------------------------------
// Self-test GEMM: D (128 x N) = A (128 x K) . B (N x K)^T from pre-packed SW128 images, one CTA.
// Validates descriptors / swizzle / TMEM mapping in isolation (tests/test_gpu_umma.py).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 1)
umma_selftest_kernel(const uint8_t *__restrict__ a_img, const uint8_t *__restrict__ b_img, int N, int K, float *__restrict__ d)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full, done;
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t a_bytes = 128u * K * 2u, b_bytes = (uint32_t)N * K * 2u;
    uint8_t *A = base, *B = base + a_bytes;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) { mbar_init(&full, 1); mbar_init(&done, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc(&tmem_base, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        mbar_arrive_expect_tx(&full, a_bytes + b_bytes);
        bulk_g2s(A, a_img, a_bytes, &full);
        bulk_g2s(B, b_img, b_bytes, &full);
        mbar_wait(&full, 0, 8);
        tc_fence_after();
        const uint32_t idesc = make_idesc_bf16(128, (uint32_t)N);
        for (int kb = 0; kb < K / 64; ++kb)
            for (int k = 0; k < 4; ++k)
                mma_bf16_ss(tmem, make_sdesc(smem_u32(A) + kb * (128 * 128) + k * 32),
                            make_sdesc(smem_u32(B) + kb * (N * 128) + k * 32), idesc, (kb | k) > 0);
        mma_commit(&done);
    }
    mbar_wait(&done, 0, 9);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int e = 0; e < 32; ++e) d[(size_t)tid * N + c0 + e] = __uint_as_float(v[e]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}


// =============================================================================================
// chain2_kernel — CTA-PAIR version (cluster of 2, tcgen05 cta_group::2).
//   * Each CTA of the pair runs the front chain for its own 128-point tile; every MMA spans both SMs
//     (M = 256): mid layers  D[256 pts, N] = A[256 pts, 64] . W^T  (each CTA holds N/2 weight rows),
//     final layer            D[256 ch, 128 pts] = W3^T[256 ch, 128] . H2^T  (each CTA holds 128 channel
//     rows of W3 and contributes 64 of the 128 point columns).
//   * W3 is RESIDENT: each CTA keeps its 4 x [128 ch x 128 K] bf16 tiles (128 KB) in shared memory for the
//     whole persistent kernel — zero steady-state weight traffic (v1 re-streamed 256 KB per tile from L2).
//   * H2 is single-buffered but split in two 64-point halves with their own full/empty barriers and the
//     final layer walks (half, channel-block): half 0 of the next tile is refilled while half 1 computes.
//   * Only the leader CTA issues MMAs; completions are multicast to both CTAs' mbarriers; the follower's
//     threads arrive on the leader's barriers through mapa + mbarrier.arrive.release.cluster.
// =============================================================================================
constexpr uint32_t kC2OffW3 = 0, kC2OffA0 = 131072, kC2OffA1 = 147456, kC2OffH2 = 163840, kC2OffWM = 196608, kC2OffWC = 212992;
constexpr uint32_t kChain2Smem = kC2OffWC + 2 * 4096 + 1024;
__device__ __forceinline__ uint32_t wm2_slot(int s) { return s == 0 ? 0u : (s == 1 ? 4096u : 8192u); }

constexpr int kTraceEvents = 2048;
constexpr uint64_t kFinalInFlight = 2;   // final-layer accumulator granules allowed in the tensor queue at once
// role: 0 front (warp 0), 1 final epilogue (warp 4), 2 MMA issuer, 3 loader; slot = (cta*4 + role)
template <bool kTrace>
__device__ __forceinline__ void trace_ev(const ChainParams &p, uint32_t rank, int role, uint32_t &cnt, uint32_t code)
{
    if (kTrace && p.trace && blockIdx.x < 2 && cnt < kTraceEvents) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[((size_t)(rank * 4 + role) * kTraceEvents + cnt) * 2] = code;
        p.trace[((size_t)(rank * 4 + role) * kTraceEvents + cnt) * 2 + 1] = t;
        ++cnt;
    }
}

struct Chain2Shared {
    uint64_t a_full[2], mid_full[2], wready, wres_full;     // [front group]
    uint64_t h2_full[2], h2_empty[2][2], acc_full[3], acc_empty[3], wc_full[2], wc_ready[2], wc_empty[2];   // h2_empty[step parity][half]
    uint32_t tmem_base;
    alignas(16) float l0[64 * 4];
    alignas(16) float mid_t[3][128];
};

// kTrace: the PCP_CHAIN_TRACE build of the kernel (role timelines of cluster 0); the production instantiation carries no
// trace code at all — its hot loops are sized for the instruction cache (section 4.1 of DESIGN.md)
template <bool kTrace>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kChainThreads, 1)
chain2_kernel(const ChainParams p)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ Chain2Shared sh;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *W3 = base + kC2OffW3, *A0 = base + kC2OffA0, *A1 = base + kC2OffA1, *H2 = base + kC2OffH2, *WM = base + kC2OffWM, *WC = base + kC2OffWC;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_ctarank();
    const int64_t ncl = gridDim.x >> 1, cl = blockIdx.x >> 1;
    const int ppc = (p.tiles_per_cloud + 1) >> 1;                        // pair-steps per cloud
    const int64_t total_steps = (int64_t)ppc * p.b;
    const int64_t s_begin = cl * total_steps / ncl, s_end = (cl + 1) * total_steps / ncl;
    const int64_t nsteps = s_end - s_begin;

    if (tid == 0) {
        for (int i = 0; i < 2; ++i) { mbar_init(&sh.a_full[i], 8); mbar_init(&sh.mid_full[i], 1); }
        mbar_init(&sh.wready, 2); mbar_init(&sh.wres_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&sh.h2_full[i], 4); mbar_init(&sh.h2_empty[0][i], 1); mbar_init(&sh.h2_empty[1][i], 1);
            mbar_init(&sh.wc_full[i], 1); mbar_init(&sh.wc_ready[i], 2); mbar_init(&sh.wc_empty[i], 1);
        }
        for (int i = 0; i < 3; ++i) { mbar_init(&sh.acc_full[i], 1); mbar_init(&sh.acc_empty[i], 8); }
        fence_barrier_init();
    }
    if (warp == 9) tmem_alloc2(&sh.tmem_base, 512);
    for (int i = tid; i < 256; i += kChainThreads) sh.l0[i] = p.l0[i];
    for (int l = 0; l < p.nmid; ++l)
        for (int i = tid; i < p.mid_n[l]; i += kChainThreads) sh.mid_t[l][i] = p.mid_t[l][i];
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = sh.tmem_base;
    // programmatic dependent launch (the launch opts in): everything above — and the loader's resident weight copies below —
    // touches only weights, so it overlaps the tail of the preceding kernel; centroid / scale / R / per-cloud weight images /
    // the pooled-feature buffer are the predecessor's output and are read or written only after this wait
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if (warp != 9) asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp < 4 || warp >= 12) {
        // =========================== front (both CTAs): L0 + mid-layer epilogues ===========================
        // TWO front groups (round 2), each with its own operand buffer (A0 / A1, written IN PLACE by every epilogue: mid_full says
        // the MMA that read the buffer is complete), its own mid accumulator in TMEM and its own barriers: group g takes the
        // pair-steps it = g, g + 2, ...  A tile still crosses seven serial hand-offs of ~650 ns, but two tiles are in flight, so
        // the final layer is fed twice as often (the traced single front chain took 5.5 us per step against 2.3 us of final MMAs).
        const int fgp = warp >= 12 ? 1 : 0, fw = warp & 3;
        const int row = fw * 32 + lane;
        const int half = fw >> 1;
        uint8_t *Ag = fgp ? A1 : A0;
        uint32_t mid_phase = 0;
        uint32_t tcnt = 0;
        const bool tr = (tid == p.trace_tid);              // (trace: one thread, by default of group 0)
        // normalised (and R-rotated) coordinates of this thread's point for pair-step `it`; evaluated one step
        // ahead so the global-load + divide latency hides behind the wait for the first mid-layer MMA
        auto load_point = [&](int64_t it, float &q0, float &q1, float &q2) {
            const int64_t st = s_begin + it;
            const int64_t cloud = st / ppc;
            const int64_t tile = 2 * (st % ppc) + rank;
            const int64_t pi = min(tile * (int64_t)kTP + row, p.n - 1);      // clamp: duplicates do not change a max
            const float *px = p.x + (cloud * p.n + pi) * 3;
            const float cx = __ldg(p.centroid + cloud * 3), cy = __ldg(p.centroid + cloud * 3 + 1), cz = __ldg(p.centroid + cloud * 3 + 2);
            const float sc = __ldg(p.scale + cloud);
            q0 = __fdiv_rn(__fsub_rn(__ldg(px), cx), sc);                    // PointNet.py:691-706
            q1 = __fdiv_rn(__fsub_rn(__ldg(px + 1), cy), sc);
            q2 = __fdiv_rn(__fsub_rn(__ldg(px + 2), cz), sc);
            if (p.R) {                                                       // X = pc @ R (PointNet.py:207)
                const float *R = p.R + cloud * 9;
                const float r0 = q0 * __ldg(R) + q1 * __ldg(R + 3) + q2 * __ldg(R + 6);
                const float r1 = q0 * __ldg(R + 1) + q1 * __ldg(R + 4) + q2 * __ldg(R + 7);
                const float r2 = q0 * __ldg(R + 2) + q1 * __ldg(R + 5) + q2 * __ldg(R + 8);
                q0 = r0; q1 = r1; q2 = r2;
            }
        };
        float q0 = 0.f, q1 = 0.f, q2 = 0.f, nq0 = 0.f, nq1 = 0.f, nq2 = 0.f;
        if (nsteps > fgp) load_point(fgp, nq0, nq1, nq2);
        for (int64_t it = fgp; it < nsteps; it += 2) {
            if (tr) trace_ev<kTrace>(p, rank, 0, tcnt, 0x100);
            q0 = nq0; q1 = nq1; q2 = nq2;
#pragma unroll 1                                           // code size: see the final epilogue below
            for (int c = 0; c < 8; ++c) {
                float o[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const float4 w = *reinterpret_cast<const float4 *>(&sh.l0[(c * 8 + e) * 4]);
                    const float v = fmaf(q2, w.z, fmaf(q1, w.y, fmaf(q0, w.x, w.w)));
                    o[e] = v > 0.f ? v : 0.f;
                }
                *reinterpret_cast<uint4 *>(Ag + sw128_offset(128, row, c * 8)) =
                    make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
            }
            // one arrival per WARP (8 per pair), not per thread: 256 serialized smem atomics per hand-off
            // were the dominant latency of the whole pipeline
            fence_proxy_async();
            __syncwarp();
            // Follower CTA -> leader barrier: relaxed cluster-scope arrive.  The payload (this CTA's smem) was made
            // visible to the async proxy by fence.proxy.async above, in program order before the arrive; a
            // release.cluster here costs ~0.5 us of MEMBAR on the critical path of every layer.
            if (lane == 0) { if (rank == 0) mbar_arrive(&sh.a_full[fgp]); else mbar_arrive_cluster_relaxed(&sh.a_full[fgp], 0); }
            if (tr) trace_ev<kTrace>(p, rank, 0, tcnt, 0x101);
            if (it + 2 < nsteps) load_point(it + 2, nq0, nq1, nq2);
            for (int l = 0; l < p.nmid; ++l) {
                mbar_wait(&sh.mid_full[fgp], mid_phase, 11);
                if (tr) trace_ev<kTrace>(p, rank, 0, tcnt, 0x110 + l);
                mid_phase ^= 1;
                tc_fence_after();
                const bool last = (l == p.nmid - 1);
                uint8_t *out = last ? H2 : Ag;
                // H2 is written by the two groups in turn: step it may be written once step it - 1 has been consumed.  One barrier per
                // step parity, so that a waiter is never more than one phase behind (a single barrier would skip phases per group
                // and a parity wait would alias): completion k of h2_empty[q] = step 2 k + q consumed.
                if (last && it >= 1) mbar_wait(&sh.h2_empty[(it - 1) & 1][half], (uint32_t)(((it - 1) >> 1) & 1), 12);
                if (tr && last) trace_ev<kTrace>(p, rank, 0, tcnt, 0x120);
                const int N = p.mid_n[l];
                for (int c0 = 0; c0 < N; c0 += 64) {
                    uint32_t va[32], vb[32];
                    const uint32_t taddr = tmem + ((uint32_t)(fw * 32) << 16) + (uint32_t)(fgp * 128) + c0;
                    tmem_ld32(taddr, va); tmem_ld32(taddr + 32, vb);
                    tmem_ld_wait();
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        float o[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const uint32_t raw = c < 4 ? va[c * 8 + e] : vb[(c - 4) * 8 + e];
                            const float f = __uint_as_float(raw) + sh.mid_t[l][c0 + c * 8 + e];
                            o[e] = f > 0.f ? f : 0.f;
                        }
                        *reinterpret_cast<uint4 *>(out + sw128_offset(128, row, c0 + c * 8)) =
                            make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
                    }
                }
                tc_fence_before();
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) {
                    uint64_t *bar = last ? &sh.h2_full[half] : &sh.a_full[fgp];
                    if (rank == 0) mbar_arrive(bar); else mbar_arrive_cluster_relaxed(bar, 0);
                }
                if (tr) trace_ev<kTrace>(p, rank, 0, tcnt, 0x130 + l);
            }
        }
    } else if (warp < 8) {
        // =========================== final epilogue (both CTAs) ===========================
        const int ch = tid - 128;
        const uint32_t lane_base = (uint32_t)((warp - 4) * 32) << 16;
        float mx[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) mx[j] = -3.0e38f;
        int64_t cur_cloud = -1;
        uint64_t g = 0;
        uint32_t tcnt = 0;
        auto flush = [&](int64_t cloud) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int c = (j * 2 + (int)rank) * 128 + ch;
                float v = mx[j] + __ldg(p.t3 + c);
                v = v > 0.f ? v : 0.f;
                atomicMax(reinterpret_cast<unsigned int *>(p.gmax + cloud * 1024 + c), __float_as_uint(v));
                mx[j] = -3.0e38f;
            }
        };
        for (int64_t it = 0; it < nsteps; ++it) {
            const int64_t cloud = (s_begin + it) / ppc;
            if (cloud != cur_cloud) {
                if (cur_cloud >= 0) flush(cur_cloud);
                cur_cloud = cloud;
            }
            // eight 128-channel blocks per step, ONE copy of the body: the four running maxima sit in a rotating register queue
            // (block hj & 3 is at its head; eight rotations are the identity, so flush() sees them in order).  Unrolled eight
            // times this role alone was 16 KB of code and the kernel's hot code overflowed the instruction cache (11 % of its
            // stall samples were stall_no_inst); re-rolling it and the L0 loop took the chain from 0.777 to 0.89 of the bf16 peak.
#pragma unroll 1
            for (int hj = 0; hj < 8; ++hj) {
                const uint32_t buf = (uint32_t)(g & 1);      // two TMEM accumulator buffers (TMEM: 2 x 128 mid + 2 x 128 final)
                mbar_wait(&sh.acc_full[buf], (uint32_t)((g >> 1) & 1), 13);
                if (tid == 128) trace_ev<kTrace>(p, rank, 1, tcnt, 0x200 + hj);
                tc_fence_after();
                // two passes of 64 columns (512 threads: 128 registers each); the buffer is released after the second load
                const uint32_t taddr = tmem + lane_base + 256 + buf * 128;
                float m0 = mx[0], m1 = -3.0e38f, m2 = -3.0e38f, m3 = -3.0e38f;      // independent chains (FMNMX3)
                {
                    uint32_t v0[32], v1[32];
                    tmem_ld32(taddr, v0); tmem_ld32(taddr + 32, v1);
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 32; e += 2) {
                        m0 = fmaxf(fmaxf(m0, __uint_as_float(v0[e])), __uint_as_float(v0[e + 1]));
                        m1 = fmaxf(fmaxf(m1, __uint_as_float(v1[e])), __uint_as_float(v1[e + 1]));
                    }
                }
                {
                    uint32_t v2[32], v3[32];
                    tmem_ld32(taddr + 64, v2); tmem_ld32(taddr + 96, v3);
                    tmem_ld_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster_relaxed(&sh.acc_empty[buf], 0);
                    if (tid == 128) trace_ev<kTrace>(p, rank, 1, tcnt, 0x210 + hj);
#pragma unroll
                    for (int e = 0; e < 32; e += 2) {
                        m2 = fmaxf(fmaxf(m2, __uint_as_float(v2[e])), __uint_as_float(v2[e + 1]));
                        m3 = fmaxf(fmaxf(m3, __uint_as_float(v3[e])), __uint_as_float(v3[e + 1]));
                    }
                }
                const float m = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
                mx[0] = mx[1]; mx[1] = mx[2]; mx[2] = mx[3]; mx[3] = m;         // rotate: block hj & 3 is at the head again after four
                ++g;
            }
        }
        if (cur_cloud >= 0) flush(cur_cloud);
    } else if (warp == 8 || warp == 11) {
        // =========================== mid-layer MMA issuer (leader CTA only) ===========================
        // Two independent issuing warps (mid layers here, final layer in warp 10) replace a single polling
        // state machine: every hand-off is a hardware-suspended mbarrier wait (wake-up ~100 cycles) instead of
        // a ~150-cycle test_wait poll per barrier per iteration, which had made the issuer the bottleneck.
        const int ig = warp == 11 ? 1 : 0;                 // one issuing warp per front group
        if (rank == 0 && nsteps > ig) {
            mbar_wait_cluster(&sh.wready, 0, 14);
            const uint32_t lo_a0 = (smem_u32(A0) & 0x3FFFFu) >> 4, lo_a1 = (smem_u32(A1) & 0x3FFFFu) >> 4;
            const uint32_t lo_wm = (smem_u32(WM) & 0x3FFFFu) >> 4, lo_wc = (smem_u32(WC) & 0x3FFFFu) >> 4;
            uint32_t a_cnt = 0, tcnt = 0;
            for (int64_t it = ig; it < nsteps; it += 2) {
                for (int l = 0; l < p.nmid; ++l) {
                    mbar_wait(&sh.a_full[ig], a_cnt & 1, 24);
                    ++a_cnt;
                    const bool percloud = p.mid_cloud_stride[l] != 0;
                    if (percloud) mbar_wait(&sh.wc_ready[it & 1], (uint32_t)((it >> 1) & 1), 25);
                    tc_fence_after();
                    const uint32_t a_lo = ig ? lo_a1 : lo_a0;
                    const uint32_t b_lo = percloud ? lo_wc + (uint32_t)(it & 1) * (4096u >> 4) : lo_wm + (wm2_slot(p.mid_slot[l]) >> 4);
                    const uint32_t idesc = make_idesc_bf16(256, (uint32_t)p.mid_n[l]);
                    if (elect_one_sync()) {
                        if (ig == 0) trace_ev<kTrace>(p, rank, 2, tcnt, 0x300 + l);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            mma2_bf16_ss(tmem + (uint32_t)(ig * 128), sdesc_from_lo(a_lo + k * 2), sdesc_from_lo(b_lo + k * 2), idesc, k > 0);
                        mma2_commit_mc(&sh.mid_full[ig], 3);
                        if (percloud) mma2_commit_mc(&sh.wc_empty[it & 1], 3);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp == 10) {
        // =========================== final-layer MMA issuer (leader CTA only) ===========================
        if (rank == 0 && nsteps > 0) {
            mbar_wait_cluster(&sh.wready, 0, 26);
            const uint32_t lo_h2 = (smem_u32(H2) & 0x3FFFFu) >> 4, lo_w3 = (smem_u32(W3) & 0x3FFFFu) >> 4;
            const uint32_t idesc_fin = make_idesc_bf16(256, 128);
            uint64_t g = 0;
            uint32_t tcnt = 0;
            for (int64_t it = 0; it < nsteps; ++it) {
                for (int hf = 0; hf < 2; ++hf) {
                    mbar_wait(&sh.h2_full[hf], (uint32_t)(it & 1), 27);
                    for (int j = 0; j < 4; ++j, ++g) {
                        const uint32_t buf = (uint32_t)(g & 1);
                        if (g >= 2) mbar_wait(&sh.acc_empty[buf], (uint32_t)(((g >> 1) - 1) & 1), 28);
                        // keep at most kFinalInFlight granules queued so a mid-layer MMA (critical path of the
                        // next tile, issued by warp 8 into the same tensor FIFO) never waits long
                        if (g >= kFinalInFlight)
                            mbar_wait(&sh.acc_full[(g - kFinalInFlight) & 1], (uint32_t)(((g - kFinalInFlight) >> 1) & 1), 29);
                        tc_fence_after();
                        const uint32_t a_lo = lo_w3 + (uint32_t)j * (32768u >> 4);
                        const uint32_t b_lo = lo_h2 + (uint32_t)hf * (8192u >> 4);
                        const uint32_t d = tmem + 256 + buf * 128;
                        if (elect_one_sync()) {
                            trace_ev<kTrace>(p, rank, 3, tcnt, 0x400 + hf * 16 + j * 2);
#pragma unroll
                            for (int kk = 0; kk < 8; ++kk)      // kk = kb*4 + k : K-block stride 16 KB, K-step 32 B
                                mma2_bf16_ss(d, sdesc_from_lo(a_lo + (kk >> 2) * (16384u >> 4) + (kk & 3) * 2),
                                             sdesc_from_lo(b_lo + (kk >> 2) * (16384u >> 4) + (kk & 3) * 2), idesc_fin, kk > 0);
                            mma2_commit_mc(&sh.acc_full[buf], 3);
                            if (j == 3) mma2_commit_mc(&sh.h2_empty[it & 1][hf], 3);
                        }
                        __syncwarp();
                    }
                }
            }
        }
    } else if (warp == 9) {
        // =========================== loader (both CTAs) ===========================
        if (lane == 0 && nsteps > 0) {
            uint32_t bytes = 8 * kW3Tile;
            for (int l = 0; l < p.nmid; ++l) if (p.mid_cloud_stride[l] == 0) bytes += (uint32_t)p.mid_n[l] * 64u;
            mbar_arrive_expect_tx(&sh.wres_full, bytes);
            for (int j = 0; j < 4; ++j)
                for (int kb = 0; kb < 2; ++kb)
                    bulk_g2s(W3 + j * 32768 + kb * 16384, p.w3 + (size_t)(((j * 2 + rank) * 2) + kb) * kW3Tile, kW3Tile, &sh.wres_full);
            for (int l = 0; l < p.nmid; ++l)
                if (p.mid_cloud_stride[l] == 0)
                    bulk_g2s(WM + wm2_slot(p.mid_slot[l]), p.mid_w[l] + (size_t)rank * p.mid_n[l] * 64u, (uint32_t)p.mid_n[l] * 64u, &sh.wres_full);
            asm volatile("griddepcontrol.wait;" ::: "memory");
            mbar_wait(&sh.wres_full, 0, 15);
            mbar_arrive_cluster(&sh.wready, 0);
            int pc_layer = -1;
            for (int l = 0; l < p.nmid; ++l) if (p.mid_cloud_stride[l] != 0) pc_layer = l;
            if (pc_layer >= 0) {
                for (int64_t it = 0; it < nsteps; ++it) {
                    const int s = (int)(it & 1);
                    if (it >= 2) mbar_wait(&sh.wc_empty[s], (uint32_t)(((it >> 1) - 1) & 1), 16);
                    const int64_t cloud = (s_begin + it) / ppc;
                    mbar_arrive_expect_tx(&sh.wc_full[s], 4096);
                    bulk_g2s(WC + s * 4096, p.mid_w[pc_layer] + cloud * p.mid_cloud_stride[pc_layer] + (size_t)rank * 4096, 4096, &sh.wc_full[s]);
                    mbar_wait(&sh.wc_full[s], (uint32_t)((it >> 1) & 1), 17);
                    mbar_arrive_cluster(&sh.wc_ready[s], 0);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 9) tmem_dealloc2(tmem, 512);
}

// 2-CTA self-test: D (256 x N) = A (256 x K) . B (N x K)^T with cta_group::2 (tests/test_gpu_umma.py)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
umma2_selftest_kernel(const uint8_t *__restrict__ a_img, const uint8_t *__restrict__ b_img, int N, int K, float *__restrict__ d)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full, ready, done;
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const uint32_t rank = cluster_ctarank();
    const uint32_t a_bytes = 128u * K * 2u, bh_bytes = (uint32_t)(N / 2) * K * 2u;
    uint8_t *A = base, *B = base + a_bytes;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) { mbar_init(&full, 1); mbar_init(&ready, 2); mbar_init(&done, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc2(&tmem_base, 256);
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    if (tid == 0) {
        // per K-block: this CTA's 128 A rows and its N/2 B rows (contiguous inside each K-block of the full images)
        mbar_arrive_expect_tx(&full, a_bytes + bh_bytes);
        for (int kb = 0; kb < K / 64; ++kb) {
            bulk_g2s(A + kb * (128 * 128), a_img + (size_t)kb * (256 * 128) + (size_t)rank * (128 * 128), 128 * 128, &full);
            bulk_g2s(B + kb * ((N / 2) * 128), b_img + (size_t)kb * (N * 128) + (size_t)rank * ((N / 2) * 128), (N / 2) * 128, &full);
        }
        mbar_wait(&full, 0, 18);
        mbar_arrive_cluster(&ready, 0);
        if (rank == 0) {
            mbar_wait_cluster(&ready, 0, 19);
            tc_fence_after();
            const uint32_t idesc = make_idesc_bf16(256, (uint32_t)N);
            for (int kb = 0; kb < K / 64; ++kb)
                for (int k = 0; k < 4; ++k)
                    mma2_bf16_ss(tmem, make_sdesc(smem_u32(A) + kb * (128 * 128) + k * 32),
                                 make_sdesc(smem_u32(B) + kb * ((N / 2) * 128) + k * 32), idesc, (kb | k) > 0);
            mma2_commit_mc(&done, 3);
        }
    }
    mbar_wait(&done, 0, 20);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int e = 0; e < 32; ++e) d[(size_t)(rank * 128 + tid) * N + c0 + e] = __uint_as_float(v[e]);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 0) tmem_dealloc2(tmem, 256);
}


// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
static inline uint16_t f2bf(float f)
{
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

// W: (cin x cout) row-major fp32; image rows = cout, K = cin (cin multiple of 64)
static void pack_b_image(const float *W, int cin, int cout, std::vector<uint16_t> &img)
{
    img.assign((size_t)cin * cout, 0);
    for (int r = 0; r < cout; ++r)
        for (int k = 0; k < cin; ++k) img[sw128_offset((uint32_t)cout, (uint32_t)r, (uint32_t)k) / 2] = f2bf(W[(size_t)k * cout + r]);
}

// final layer 128 -> 1024: 16 tiles (j, h): rows = channels j*128.., K = inputs h*64..
static void pack_w3_image(const float *W, std::vector<uint16_t> &img)
{
    img.assign((size_t)128 * 1024, 0);
    for (int j = 0; j < 8; ++j)
        for (int h = 0; h < 2; ++h) {
            uint16_t *tile = img.data() + (size_t)(j * 2 + h) * (kW3Tile / 2);
            for (int r = 0; r < 128; ++r)
                for (int k = 0; k < 64; ++k) tile[sw128_offset(128, (uint32_t)r, (uint32_t)k) / 2] = f2bf(W[(size_t)(h * 64 + k) * 1024 + j * 128 + r]);
        }
}

struct TcChain {
    uint8_t *mid_w[3] = {nullptr, nullptr, nullptr};
    const float *mid_t[3] = {nullptr, nullptr, nullptr};
    int mid_n[3] = {0, 0, 0};
    int nmid = 0;
    uint8_t *w3 = nullptr;
    const float *t3 = nullptr;
    float *l0 = nullptr;
};

struct TcState {
    TcChain A, B, C;
    uint8_t *wc = nullptr; size_t wc_clouds = 0;     // per-cloud folded (R64 . W_s2_1) images
    float *centroid = nullptr, *scale = nullptr; size_t stat_clouds = 0;
    std::vector<void *> allocs;
    bool attr_set = false;
    bool timing = false;                                  // CUDA-event pairs around every chain launch (bench.py roofline)
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> events;
};

static int dev_upload(TcState *ts, const void *src, size_t bytes, void **dst)
{
    PCP_CUDA_OK(cudaMalloc(dst, bytes));
    ts->allocs.push_back(*dst);
    PCP_CUDA_OK(cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
    return PCP_OK;
}

static int build_chain(pcp_model *m, TcState *ts, TcChain &c, int l0_id, const int *mid_ids, int nmid, int fin_id,
                       const std::vector<float> &fold)
{
    const LayerSpec &L0 = m->layers[l0_id];
    std::vector<float> l0(256);
    for (int ch = 0; ch < 64; ++ch) {
        for (int k = 0; k < 3; ++k) l0[ch * 4 + k] = fold[L0.fold_w + (size_t)k * 64 + ch];
        l0[ch * 4 + 3] = fold[L0.fold_t + ch];
    }
    PCP_TRY(dev_upload(ts, l0.data(), l0.size() * 4, reinterpret_cast<void **>(&c.l0)));
    c.nmid = nmid;
    std::vector<uint16_t> img;
    for (int i = 0; i < nmid; ++i) {
        if (mid_ids[i] < 0) { c.mid_w[i] = nullptr; c.mid_n[i] = 64; c.mid_t[i] = nullptr; continue; }
        const LayerSpec &L = m->layers[mid_ids[i]];
        pack_b_image(fold.data() + L.fold_w, L.cin, L.cout, img);
        PCP_TRY(dev_upload(ts, img.data(), img.size() * 2, reinterpret_cast<void **>(&c.mid_w[i])));
        c.mid_n[i] = L.cout;
        c.mid_t[i] = m->d_fold + L.fold_t;
    }
    const LayerSpec &F = m->layers[fin_id];
    pack_w3_image(fold.data() + F.fold_w, img);
    PCP_TRY(dev_upload(ts, img.data(), img.size() * 2, reinterpret_cast<void **>(&c.w3)));
    c.t3 = m->d_fold + F.fold_t;
    return PCP_OK;
}

void model_destroy_bf16(pcp_model *m)
{
    heads_tc_destroy(m);
    if (!m->tc) return;
    TcState *ts = reinterpret_cast<TcState *>(m->tc);
    for (void *p : ts->allocs) PCP_RELEASE(cudaFree(p));
    PCP_RELEASE(cudaFree(ts->wc)); PCP_RELEASE(cudaFree(ts->centroid)); PCP_RELEASE(cudaFree(ts->scale));
    delete ts;
    m->tc = nullptr;
}

int model_prepare_bf16(pcp_model *m)
{
    model_destroy_bf16(m);
    TcState *ts = new TcState();
    m->tc = ts;
    std::vector<float> fold(m->n_fold);
    PCP_CUDA_OK(cudaMemcpy(fold.data(), m->d_fold, m->n_fold * sizeof(float), cudaMemcpyDeviceToHost));
    PCP_TRY(heads_tc_prepare(m, fold));
    if (!m->desc.vanilla) {
        const int midA[1] = {L_IT_C2};
        PCP_TRY(build_chain(m, ts, ts->A, L_IT_C1, midA, 1, L_IT_C3, fold));
        const int midB[3] = {L_S1_2, L_FT_C1, L_FT_C2};
        PCP_TRY(build_chain(m, ts, ts->B, L_S1_1, midB, 3, L_FT_C3, fold));
        const int midC[3] = {L_S1_2, -1, L_S2_2};                       // -1: per-cloud R64 . W(s2_l1_64)
        PCP_TRY(build_chain(m, ts, ts->C, L_S1_1, midC, 3, L_S2_3, fold));
        ts->C.mid_t[1] = m->d_fold + m->layers[L_S2_1].fold_t;
    } else {
        const int midC[3] = {L_S1_2, L_S2_1, L_S2_2};
        PCP_TRY(build_chain(m, ts, ts->C, L_S1_1, midC, 3, L_S2_3, fold));
    }
    return PCP_OK;
}

// The per-cloud operand images image_b(n, k) = bf16( sum_m R64_b[k][m] * W[m][n] ) — X_64 . W == X . (R64 . W)
// (PointNet.py:228,236) — come straight out of the feature T-Net's last GEMM (heads_tc.cu: heads_tc_tnet_tail_fold).

// d_x / gmax (/ R / wc) point at the first cloud of the launch; cloud0 is its index in the batch the statistics were taken over
static int launch_chain(TcState *ts, const TcChain &c, const float *d_x, int b, int64_t n, const float *R, const uint8_t *wc,
                        float *gmax, cudaStream_t st, int cloud0 = 0)
{
    ChainParams p;
    memset(&p, 0, sizeof(p));
    p.x = d_x; p.n = n; p.b = b; p.centroid = ts->centroid + (size_t)cloud0 * 3; p.scale = ts->scale + cloud0; p.R = R; p.l0 = c.l0;
    p.nmid = c.nmid;
    for (int i = 0; i < c.nmid; ++i) {
        p.mid_n[i] = c.mid_n[i];
        p.mid_t[i] = c.mid_t[i];
        p.mid_w[i] = c.mid_w[i];
        p.mid_cloud_stride[i] = 0;
        if (!c.mid_w[i]) { p.mid_w[i] = wc; p.mid_cloud_stride[i] = 8192; }
        // slot: 8 KB slots 0/1 for 64-row images, the 16 KB slot 2 for the 128-row image
        p.mid_slot[i] = (c.mid_n[i] == 128) ? 2 : (i == 0 ? 0 : 1);
    }
    p.w3 = c.w3; p.t3 = c.t3; p.gmax = gmax;
    p.tiles_per_cloud = (int)((n + kTP - 1) / kTP);
    p.total_tiles = (int64_t)p.tiles_per_cloud * b;
    if (!ts->attr_set) {
        PCP_CUDA_OK(cudaFuncSetAttribute(chain2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kChain2Smem));
        PCP_CUDA_OK(cudaFuncSetAttribute(chain2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kChain2Smem));
        ts->attr_set = true;
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ts->timing) {
        PCP_CUDA_OK(cudaEventCreate(&e0)); PCP_CUDA_OK(cudaEventCreate(&e1));
        PCP_CUDA_OK(cudaEventRecord(e0, st));
    }
    const int64_t steps = (int64_t)((p.tiles_per_cloud + 1) / 2) * b;
    const int ncl = (int)std::min<int64_t>(steps, sm_count() / 2);
    static const char *trace_path = getenv("PCP_CHAIN_TRACE");          // development aid: dump role timelines of cluster 0
    unsigned long long *d_trace = nullptr;
    const size_t trace_bytes = (size_t)8 * kTraceEvents * 2 * sizeof(unsigned long long);
    if (trace_path) {
        PCP_CUDA_OK(cudaMalloc(&d_trace, trace_bytes));
        PCP_CUDA_OK(cudaMemsetAsync(d_trace, 0, trace_bytes, st));
        p.trace = d_trace;
        if (const char *e = getenv("PCP_CHAIN_TRACE_TID")) p.trace_tid = atoi(e);
    }
    {
        static const bool no_pdl = getenv("PCP_TC_GEMM_NO_PDL") != nullptr;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2 * ncl); cfg.blockDim = dim3(kChainThreads); cfg.dynamicSmemBytes = kChain2Smem; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = no_pdl ? 0 : 1;
        if (trace_path) PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, chain2_kernel<true>, p));
        else PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, chain2_kernel<false>, p));
    }
    if (trace_path) {
        std::vector<unsigned long long> h(trace_bytes / 8);
        PCP_CUDA_OK(cudaStreamSynchronize(st));
        PCP_CUDA_OK(cudaMemcpy(h.data(), d_trace, trace_bytes, cudaMemcpyDeviceToHost));
        cudaFree(d_trace);
        static int seq = 0;
        char name[512];
        snprintf(name, sizeof(name), "%s.%d.bin", trace_path, seq++);
        if (FILE *f = fopen(name, "wb")) { fwrite(h.data(), 1, trace_bytes, f); fclose(f); }
    }
    PCP_CHECK_LAUNCH();
    if (ts->timing) {
        PCP_CUDA_OK(cudaEventRecord(e1, st));
        ts->events.emplace_back(e0, e1);
    }
    return PCP_OK;
}

// The bf16 forward in three parts, so that a caller whose points are still arriving from the host can run the part that only
// needs the points (statistics + the first chain) on each chunk as it lands and everything behind it once, on the whole batch:
//   begin(b)              buffers for b clouds, pooled features cleared
//   stage1(x, c0, bc)     clouds [c0, c0 + bc): centroid / scale, then chain A (full net) or the whole trunk (vanilla) -> pooled rows
//   finish(x, b, ...)     T-Net tails, chains B and C, heads
int model_forward_bf16_begin(pcp_model *m, int b, cudaStream_t st)
{
    TcState *ts = reinterpret_cast<TcState *>(m->tc);
    PCP_REQUIRE(ts, PCP_ERR_STATE, "bf16 path: model not prepared");
    Workspace &ws = m->ws;
    if (ts->stat_clouds < (size_t)b) {
        cudaFree(ts->centroid); cudaFree(ts->scale); cudaFree(ts->wc);
        ts->centroid = ts->scale = nullptr; ts->wc = nullptr; ts->stat_clouds = 0;
        PCP_CUDA_OK(cudaMalloc(&ts->centroid, (size_t)b * 3 * sizeof(float)));
        PCP_CUDA_OK(cudaMalloc(&ts->scale, (size_t)b * sizeof(float)));
        PCP_CUDA_OK(cudaMalloc(&ts->wc, (size_t)b * 8192));
        ts->stat_clouds = (size_t)b;
    }
    PCP_CUDA_OK(cudaMemsetAsync(ws.g3, 0, (size_t)b * 1024 * sizeof(float), st));
    if (!m->desc.vanilla) {
        PCP_CUDA_OK(cudaMemsetAsync(ws.g1, 0, (size_t)b * 1024 * sizeof(float), st));
        PCP_CUDA_OK(cudaMemsetAsync(ws.g2, 0, (size_t)b * 1024 * sizeof(float), st));
    }
    return PCP_OK;
}

int model_forward_bf16_stage1(pcp_model *m, const float *d_x_chunk, int c0, int bc, int64_t n, cudaStream_t st)
{
    TcState *ts = reinterpret_cast<TcState *>(m->tc);
    Workspace &ws = m->ws;
    // per-cloud centroid / scale only (the chains normalise on the fly; nothing per-point is written)
    PCP_TRY(pcp_normalize_f32(d_x_chunk, bc, n, nullptr, ts->centroid + (size_t)c0 * 3, ts->scale + c0, st));
    if (!m->desc.vanilla) return launch_chain(ts, ts->A, d_x_chunk, bc, n, nullptr, nullptr, ws.g1 + (size_t)c0 * 1024, st, c0);
    return launch_chain(ts, ts->C, d_x_chunk, bc, n, nullptr, nullptr, ws.g3 + (size_t)c0 * 1024, st, c0);
}

int model_forward_bf16_finish(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg, float *d_R,
                              float *d_cls_logits, float *d_seg_logits, cudaStream_t st)
{
    TcState *ts = reinterpret_cast<TcState *>(m->tc);
    Workspace &ws = m->ws;
    const bool full = !m->desc.vanilla;
    float *Rout = d_R ? d_R : ws.R;
    if (full) {
        PCP_TRY(heads_tc_tnet_tail(m, 0, ws.g1, b, Rout, st));
        PCP_TRY(launch_chain(ts, ts->B, d_x, b, n, Rout, nullptr, ws.g2, st));
        // the seg head folds R64 into its own first layer; pcp_model_evaluate wants its orthogonality regulariser
        const bool need_r64 = ((heads & PCP_HEAD_SEG) && (d_seg || d_seg_logits)) || m->eval_reg_acc;
        PCP_TRY(heads_tc_tnet_tail_fold(m, ws.g2, b, ts->wc, need_r64 ? ws.R64 : nullptr, st));
        if (m->eval_reg_acc) PCP_TRY(tnet_reg_accumulate(ws.R64, b, 64, m->eval_reg_acc, st));
        PCP_TRY(launch_chain(ts, ts->C, d_x, b, n, Rout, ts->wc, ws.g3, st));
    } else if (d_R) {
        PCP_TRY(model_eye3(d_R, b, st));
    }
    if ((heads & PCP_HEAD_CLS) && (d_cls || d_cls_logits)) PCP_TRY(heads_tc_cls(m, ws.g3, b, d_cls, d_cls_logits, st));
    if ((heads & PCP_HEAD_SEG) && (d_seg || d_seg_logits))
        PCP_TRY(heads_tc_seg(m, d_x, b, n, ts->centroid, ts->scale, full ? Rout : nullptr, full ? ws.R64 : nullptr, ws.g3, d_seg, d_seg_logits, st));
    return PCP_OK;
}

int model_forward_bf16(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg,
                       float *d_R, float *d_cls_logits, float *d_seg_logits, cudaStream_t st)
{
    PCP_TRY(model_forward_bf16_begin(m, b, st));
    PCP_TRY(model_forward_bf16_stage1(m, d_x, 0, b, n, st));
    return model_forward_bf16_finish(m, d_x, b, n, heads, d_cls, d_seg, d_R, d_cls_logits, d_seg_logits, st);
}

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_model_enable_timing(pcp_model *m, int flag)
{
    PCP_REQUIRE(m, PCP_ERR_INVALID, "pcp_model_enable_timing: null model");
    PCP_REQUIRE(m->tc, PCP_ERR_STATE, "pcp_model_enable_timing: bf16 model not prepared");
    reinterpret_cast<TcState *>(m->tc)->timing = flag != 0;
    return PCP_OK;
}

extern "C" int pcp_model_chain_time(pcp_model *m, float *ms_total, int *launches)
{
    PCP_REQUIRE(m && ms_total && launches, PCP_ERR_INVALID, "pcp_model_chain_time: null argument");
    PCP_REQUIRE(m->tc, PCP_ERR_STATE, "pcp_model_chain_time: bf16 model not prepared");
    TcState *ts = reinterpret_cast<TcState *>(m->tc);
    float tot = 0.f;
    for (auto &ev : ts->events) {
        PCP_CUDA_OK(cudaEventSynchronize(ev.second));
        float ms = 0.f;
        PCP_CUDA_OK(cudaEventElapsedTime(&ms, ev.first, ev.second));
        tot += ms;
        cudaEventDestroy(ev.first); cudaEventDestroy(ev.second);
    }
    *ms_total = tot;
    *launches = (int)ts->events.size();
    ts->events.clear();
    return PCP_OK;
}

// D (M x N) = A (M x K) . B (N x K)^T through tcgen05 (host fp32 in/out; operands rounded to bf16).
// M = 128: one CTA, cta_group::1.  M = 256: CTA pair, cta_group::2.
static int selftest_umma_pair(const float *h_a, const float *h_b, int N, int K, float *h_d)
{
    std::vector<uint16_t> ai((size_t)256 * K), bi((size_t)N * K);
    for (int r = 0; r < 256; ++r) for (int k = 0; k < K; ++k) ai[sw128_offset(256, r, k) / 2] = f2bf(h_a[(size_t)r * K + k]);
    for (int r = 0; r < N; ++r) for (int k = 0; k < K; ++k) bi[sw128_offset(N, r, k) / 2] = f2bf(h_b[(size_t)r * K + k]);
    uint8_t *da = nullptr, *db = nullptr; float *dd = nullptr;
    PCP_CUDA_OK(cudaMalloc(&da, ai.size() * 2)); PCP_CUDA_OK(cudaMalloc(&db, bi.size() * 2)); PCP_CUDA_OK(cudaMalloc(&dd, (size_t)256 * N * 4));
    PCP_CUDA_OK(cudaMemcpy(da, ai.data(), ai.size() * 2, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(db, bi.data(), bi.size() * 2, cudaMemcpyHostToDevice));
    const size_t smem = (size_t)128 * K * 2 + (size_t)(N / 2) * K * 2 + 1024;
    PCP_CUDA_OK(cudaFuncSetAttribute(umma2_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma2_selftest_kernel<<<2, 128, smem>>>(da, db, N, K, dd);
    PCP_CHECK_LAUNCH();
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        unsigned info[4] = {0, 0, 0, 0};
        cudaMemcpyFromSymbol(info, g_mbar_timeout_info, sizeof(info));
        set_error("pcp_selftest_umma(pair): %s (timeout info %08x %u %u %u)", cudaGetErrorString(e), info[0], info[1], info[2], info[3]);
        return PCP_ERR_CUDA;
    }
    PCP_CUDA_OK(cudaMemcpy(h_d, dd, (size_t)256 * N * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(db); cudaFree(dd);
    return PCP_OK;
}

extern "C" int pcp_selftest_umma2(const float *h_a, const float *h_b, int N, int K, float *h_d)
{
    PCP_REQUIRE(h_a && h_b && h_d && N % 32 == 0 && N >= 32 && N <= 256 && K % 64 == 0 && K >= 64 && K <= 256, PCP_ERR_INVALID,
                "pcp_selftest_umma2: N in {32..256 step 32}, K in {64,128,192,256}");
    return selftest_umma_pair(h_a, h_b, N, K, h_d);
}

extern "C" int pcp_selftest_umma(const float *h_a, const float *h_b, int N, int K, float *h_d)
{
    PCP_REQUIRE(h_a && h_b && h_d && N % 32 == 0 && N >= 32 && N <= 256 && K % 64 == 0 && K >= 64 && K <= 256, PCP_ERR_INVALID,
                "pcp_selftest_umma: N in {32..256 step 32}, K in {64,128,192,256}");
    std::vector<uint16_t> ai((size_t)128 * K), bi((size_t)N * K);
    for (int r = 0; r < 128; ++r) for (int k = 0; k < K; ++k) ai[sw128_offset(128, r, k) / 2] = f2bf(h_a[(size_t)r * K + k]);
    for (int r = 0; r < N; ++r) for (int k = 0; k < K; ++k) bi[sw128_offset(N, r, k) / 2] = f2bf(h_b[(size_t)r * K + k]);
    uint8_t *da = nullptr, *db = nullptr; float *dd = nullptr;
    PCP_CUDA_OK(cudaMalloc(&da, ai.size() * 2)); PCP_CUDA_OK(cudaMalloc(&db, bi.size() * 2)); PCP_CUDA_OK(cudaMalloc(&dd, (size_t)128 * N * 4));
    PCP_CUDA_OK(cudaMemcpy(da, ai.data(), ai.size() * 2, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(db, bi.data(), bi.size() * 2, cudaMemcpyHostToDevice));
    const size_t smem = ai.size() * 2 + bi.size() * 2 + 1024;
    PCP_CUDA_OK(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma_selftest_kernel<<<1, 128, smem>>>(da, db, N, K, dd);
    PCP_CHECK_LAUNCH();
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        unsigned info[4] = {0, 0, 0, 0};
        cudaMemcpyFromSymbol(info, g_mbar_timeout_info, sizeof(info));
        set_error("pcp_selftest_umma: %s (timeout info %08x %u %u %u)", cudaGetErrorString(e), info[0], info[1], info[2], info[3]);
        return PCP_ERR_CUDA;
    }
    PCP_CUDA_OK(cudaMemcpy(h_d, dd, (size_t)128 * N * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(db); cudaFree(dd);
    return PCP_OK;
}
