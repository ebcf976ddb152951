// sampling.cu — farthest-point sampling (north_star operator; no reference implementation, the
// contract is oracle/sampling.py:fps — SURVEY.md §8 a9).
//
// Design: FPS is a serial chain of S global arg-max steps, so the working set must never leave the
// chip.  One thread-block CLUSTER per cloud; every thread keeps its points (x,y,z) and their running
// min-distance in REGISTERS for the whole kernel — or, when the batch needs more than one wave of clusters, the larger part of
// its slice's coordinates in shared memory with the distances still in registers (hybrid residency) — so the cloud is read
// from HBM exactly once.  Per iteration: distance update (3 sub, 3 mul, 2 add, min — no FMA, bit-exact with the oracle),
// REDUX warp arg-max, one __syncthreads, then the warp that owns the block's winner pushes one 32-byte record into every CTA
// of the cluster as st.async with the arrival attached (mbarrier complete_tx); every CTA waits on its own transaction
// barrier — no hardware cluster barrier in the loop.  Ties resolve to the lowest index (np.argmax).
// Exclusive use of the device is assumed: CTAs of a cluster wait for their peers' records (bounded: a lost arrival traps).
#include <cooperative_groups.h>

#include <cstdlib>

#include "common.cuh"
#include "umma.cuh"

namespace cg = cooperative_groups;

namespace pcp {

constexpr int kFpsThreads = 512;
constexpr int kFpsWarps = kFpsThreads / 32;
constexpr int kFpsMaxCluster = 16;
constexpr int kFpsRegRows = 17, kFpsRegDist = 32;  // hybrid split: rows fully in registers / rows with only the distance there
constexpr int kFpsMaxSmemRows = 27;               // 27 * 512 points * 16 B = 221 184 B of dynamic shared memory

struct __align__(16) FpsRec {
    float x, y, z, pad;       // one 16-byte read for the winner's coordinates
    unsigned long long key;   // (float bits of min-distance << 32) | (0xFFFFFFFF - point index)
    unsigned long long pad2;
};

__device__ __forceinline__ float sqdist_nofma(float px, float py, float pz, float cx, float cy, float cz) {
    const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy), dz = __fsub_rn(pz, cz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// Cluster barrier halves.  Only the warp that pushed records has anything to release; a release fence in the other fifteen
// warps of every CTA costs ~1000 cycles per iteration for nothing.
__device__ __forceinline__ void fps_arrive(bool pushed)
{
    if (pushed) asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    else asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
__device__ __forceinline__ void fps_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// 16 bytes into CTA `cta`'s copy of a shared location, completing 16 bytes of the transaction count of its copy of `bar`
__device__ __forceinline__ void fps_st_async16(const void *local_dst, const uint64_t *local_bar, uint32_t cta, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("{\n\t.reg .b32 ra, rb;\n\tmapa.shared::cluster.u32 ra, %0, %2;\n\tmapa.shared::cluster.u32 rb, %1, %2;\n\t"
                 "st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.b32 [ra], {%3, %4, %5, %6}, [rb];\n\t}"
                 ::"r"(umma::smem_u32(local_dst)), "r"(umma::smem_u32(local_bar)), "r"(cta), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

struct FpsSide {                       // block-level exchange state of a cloud
    uint64_t *xbar;                    // [2] one transaction barrier per slot parity (nullptr: hardware cluster barrier)
    unsigned long long *warp_key;      // [kFpsWarps]
    FpsRec (*slots)[kFpsMaxCluster];   // [2][kFpsMaxCluster]
};

// after the thread's (best, cand) is known: warp keys -> (caller syncs the block) -> push of the block's winner
__device__ __forceinline__ void fps_warp_key(const FpsSide &S, unsigned bb, bool real, unsigned gi, unsigned &cand, int lane, int warp)
{
    const unsigned wmax = __reduce_max_sync(0xffffffffu, bb);
    cand = (bb == wmax && real) ? gi : 0xFFFFFFFFu;
    const unsigned wmin = __reduce_min_sync(0xffffffffu, cand);
    if (lane == 0) S.warp_key[warp] = ((unsigned long long)wmax << 32) | (unsigned long long)(0xFFFFFFFFu - wmin);
}

template <typename Coord>
__device__ __forceinline__ bool fps_push(const FpsSide &S, cg::cluster_group &cluster, int par, int rank, int cs, unsigned cand, int tid, int lane,
                                         Coord &&coord)
{
    const unsigned long long key = lane < kFpsWarps ? S.warp_key[lane] : 0ull;
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
    const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
    const bool mine = mlo != 0u && cand == 0xFFFFFFFFu - mlo;
    // the warp that holds the winner (warp 0 when the slice has no real point) pushes: lane t writes CTA t's slot
    const unsigned owner = __ballot_sync(0xffffffffu, mine || (mlo == 0u && tid == 0));
    if (owner) {
        FpsRec r;
        r.key = ((unsigned long long)mhi << 32) | mlo;
        r.x = r.y = r.z = r.pad = 0.f;
        r.pad2 = 0ull;
        if (mine) coord(r);
        const int src = __ffs(owner) - 1;
        r.x = __shfl_sync(0xffffffffu, r.x, src); r.y = __shfl_sync(0xffffffffu, r.y, src); r.z = __shfl_sync(0xffffffffu, r.z, src);
        if (cs == 1) { if (lane == 0) S.slots[par][0] = r; }
        else if (lane < cs) {
            if (S.xbar) {
                // data and its arrival in one remote operation: the receiver's barrier completes when all cs records are in
                FpsRec *dst = &S.slots[par][rank];
                fps_st_async16(&dst->x, &S.xbar[par], (uint32_t)lane, __float_as_uint(r.x), __float_as_uint(r.y), __float_as_uint(r.z), 0u);
                fps_st_async16(&dst->key, &S.xbar[par], (uint32_t)lane, mlo, mhi, 0u, 0u);
            } else {
                *cluster.map_shared_rank(&S.slots[par][rank], lane) = r;
            }
        }
    }
    return owner != 0u;
}

__device__ __forceinline__ unsigned fps_winner(const FpsSide &S, int par, int cs, int lane, float &cx, float &cy, float &cz)
{
    const unsigned long long key = lane < cs ? S.slots[par][lane].key : 0ull;
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
    const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
    const int br = __ffs(__ballot_sync(0xffffffffu, lane < cs && hi == mhi && lo == mlo)) - 1;
    const float4 w = *reinterpret_cast<const float4 *>(&S.slots[par][br].x);
    cx = w.x; cy = w.y; cz = w.z;
    return 0xFFFFFFFFu - mlo;
}

// phase cycle totals of block 0 / thread 0 (development aid, read through pcp_fps_phase_cycles)
__device__ unsigned long long g_fps_phase[8];
#define FPS_STAMP(i) do { if (kProf && blockIdx.x == 0 && tid == 0) { const long long t_ = clock64(); ph[i] += t_ - tprev; tprev = t_; } } while (0)

template <int PPT, bool kProf, int SD = 0>
__global__ void __launch_bounds__(kFpsThreads, 1)
fps_cluster_kernel(const float *__restrict__ x, int64_t n, int s, int start, int32_t *__restrict__ idx_out, int cs, int64_t ppc, int hw_barrier,
                   int srows)
{   // srows: rows of 512 points the thread block keeps in SHARED memory on top of the PPT rows in registers (0: none)
    // SD == 0: shared rows hold x | y | z | d.  SD > 0: shared rows hold the coordinates only and their running distances stay
    // in SD registers per thread (srows <= SD, rows allocated in groups of four) — 12 B instead of 20 B of shared-memory
    // traffic per point and iteration.
    extern __shared__ __align__(16) float fps_dyn[];                     // SoA, arows * 512 floats each
    const int arows = SD > 0 ? (srows + 3) & ~3 : srows;
    __shared__ unsigned long long warp_key[kFpsWarps];
    __shared__ FpsRec slots[2][kFpsMaxCluster];

    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int64_t cloud = blockIdx.x / cs;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *xc = x + cloud * n * 3;
    const int64_t p0 = (int64_t)rank * ppc;
    const int64_t lim = min(n, p0 + ppc);

    float px[PPT], py[PPT], pz[PPT], d[PPT];
#pragma unroll
    for (int k = 0; k < PPT; ++k) {
        const int64_t gi = p0 + (int64_t)k * kFpsThreads + tid;
        if (gi < lim) {
            px[k] = __ldg(xc + 3 * gi); py[k] = __ldg(xc + 3 * gi + 1); pz[k] = __ldg(xc + 3 * gi + 2);
            d[k] = __int_as_float(0x7f800000);
        } else {
            px[k] = py[k] = pz[k] = 0.f;
            d[k] = 0.f;      // padding never wins: ties go to the lowest (real) index
        }
    }
    float *sx = fps_dyn, *sy = sx + arows * kFpsThreads, *sz = sy + arows * kFpsThreads, *sd = sz + arows * kFpsThreads;
    float ds[SD > 0 ? SD : 1];
    for (int k = 0; k < arows; ++k) {
        const int64_t gi = p0 + (int64_t)(PPT + k) * kFpsThreads + tid;
        const bool real = gi < lim;
        const int i = k * kFpsThreads + tid;
        sx[i] = real ? __ldg(xc + 3 * gi) : 0.f; sy[i] = real ? __ldg(xc + 3 * gi + 1) : 0.f; sz[i] = real ? __ldg(xc + 3 * gi + 2) : 0.f;
        if (SD == 0) sd[i] = real ? __int_as_float(0x7f800000) : 0.f;
    }
    if (SD > 0) {
#pragma unroll
        for (int k = 0; k < SD; ++k) ds[k] = (p0 + (int64_t)(PPT + k) * kFpsThreads + tid < lim) ? __int_as_float(0x7f800000) : 0.f;
    }
    float cx = __ldg(xc + 3 * (int64_t)start), cy = __ldg(xc + 3 * (int64_t)start + 1), cz = __ldg(xc + 3 * (int64_t)start + 2);
    if (rank == 0 && tid == 0) idx_out[cloud * s] = start;

    // exchange: records travel as st.async with the arrival attached (one DSMEM trip + a local mbarrier wake-up) instead of
    // remote stores + the hardware cluster barrier (~1300 cycles per iteration, profiles/r01_fps_phases.txt)
    __shared__ uint64_t xbar[2];
    const bool use_xbar = cs > 1 && !hw_barrier;
    const FpsSide SS{use_xbar ? xbar : nullptr, warp_key, slots};
    uint32_t xph[2] = {0u, 0u};
    if (use_xbar) {
        if (tid == 0) { umma::mbar_init(&xbar[0], 1); umma::mbar_init(&xbar[1], 1); umma::fence_barrier_init(); }
        __syncthreads();
        if (tid == 0) { umma::mbar_arrive_expect_tx(&xbar[0], cs * 32u); umma::mbar_arrive_expect_tx(&xbar[1], cs * 32u); }
        cluster.sync();                                     // every CTA is armed before anybody pushes
    }
    long long ph[6] = {0, 0, 0, 0, 0, 0}, tprev = clock64();
    for (int it = 1; it < s; ++it) {
        // distance update with the thread's arg-max carried along (strict > keeps the lowest k = lowest index; padding sits at
        // the highest k with d = 0, so it only "wins" a thread that holds no real point)
        float best = -1.f;
        int bk = 0;
#pragma unroll
        for (int k = 0; k < PPT; ++k) {
            const float dd = sqdist_nofma(px[k], py[k], pz[k], cx, cy, cz);
            d[k] = fminf(d[k], dd);
            if (d[k] > best) { best = d[k]; bk = k; }
        }
        // the rows in shared memory (each thread only ever touches its own column: no barrier); conflict-free 4-byte accesses
        if (SD > 0) {
#pragma unroll
            for (int k4 = 0; k4 < SD; k4 += 4) {
                if (k4 < arows) {
#pragma unroll
                    for (int k = k4; k < k4 + 4 && k < SD; ++k) {
                        const int i = k * kFpsThreads + tid;
                        const float m = fminf(ds[k], sqdist_nofma(sx[i], sy[i], sz[i], cx, cy, cz));
                        ds[k] = m;
                        if (m > best) { best = m; bk = PPT + k; }
                    }
                }
            }
        } else {
#pragma unroll 4
            for (int k = 0; k < srows; ++k) {
                const int i = k * kFpsThreads + tid;
                const float m = fminf(sd[i], sqdist_nofma(sx[i], sy[i], sz[i], cx, cy, cz));
                sd[i] = m;
                if (m > best) { best = m; bk = PPT + k; }
            }
        }
        FPS_STAMP(0);
        const unsigned bb = __float_as_uint(best);
        const unsigned wmax = __reduce_max_sync(0xffffffffu, bb);
        const int64_t gi = p0 + (int64_t)bk * kFpsThreads + tid;
        const unsigned cand = (bb == wmax && gi < lim) ? (unsigned)gi : 0xFFFFFFFFu;
        const unsigned wmin = __reduce_min_sync(0xffffffffu, cand);
        if (lane == 0) warp_key[warp] = ((unsigned long long)wmax << 32) | (unsigned long long)(0xFFFFFFFFu - wmin);   // low word 0: no real point
        FPS_STAMP(1);
        __syncthreads();
        FPS_STAMP(2);
        const int par = it & 1;
        // every warp reduces the 16 warp keys itself (no second barrier); the warp that owns the block's winner fetches its
        // coordinates and pushes the record into every CTA of the cluster (lane t -> CTA t)
        const bool pushed = fps_push(SS, cluster, par, rank, cs, cand, tid, lane, [&](FpsRec &r) {
            if (bk >= PPT) {
                const int i = (bk - PPT) * kFpsThreads + tid;
                r.x = sx[i]; r.y = sy[i]; r.z = sz[i];
            } else {
#pragma unroll
                for (int k = 0; k < PPT; ++k)
                    if (k == bk) { r.x = px[k]; r.y = py[k]; r.z = pz[k]; }
            }
        });
        FPS_STAMP(3);
        if (cs == 1) __syncthreads();
        else if (use_xbar) {
            umma::mbar_wait_cluster(&xbar[par], xph[par], 72);
            xph[par] ^= 1u;
            // re-arm for iteration it + 2: no peer can push that early (it needs our it + 1 push, which follows this)
            if (tid == 0 && it + 2 < s) umma::mbar_arrive_expect_tx(&xbar[par], cs * 32u);
        } else { fps_arrive(pushed); fps_wait(); }
        FPS_STAMP(4);
        {
            const unsigned long long key = lane < cs ? slots[par][lane].key : 0ull;
            const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
            const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
            const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
            const int br = __ffs(__ballot_sync(0xffffffffu, lane < cs && hi == mhi && lo == mlo)) - 1;
            const float4 w = *reinterpret_cast<const float4 *>(&slots[par][br].x);
            cx = w.x; cy = w.y; cz = w.z;
            if (rank == 0 && tid == 0) idx_out[cloud * s + it] = (int32_t)(0xFFFFFFFFu - mlo);
        }
        FPS_STAMP(5);
    }
    if (kProf && blockIdx.x == 0 && tid == 0)
        for (int i = 0; i < 6; ++i) g_fps_phase[i] = (unsigned long long)ph[i];
    if (cs > 1) cluster.sync();    // no CTA may exit while a peer can still write into its slots
}

// Fallback for clouds too large to stay register-resident: one CTA per cloud, running distances in
// a global scratch array (L2-resident for realistic sizes).  Same arithmetic and tie-break.
__global__ void __launch_bounds__(1024, 1)
fps_stream_kernel(const float *__restrict__ x, int64_t n, int s, int start, int32_t *__restrict__ idx_out, float *__restrict__ dist)
{
    __shared__ unsigned long long wkey[32];
    __shared__ unsigned long long bkey;
    const int64_t cloud = blockIdx.x;
    const float *xc = x + cloud * n * 3;
    float *dc = dist + cloud * n;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int64_t i = tid; i < n; i += 1024) dc[i] = __int_as_float(0x7f800000);
    int64_t cur = start;
    if (tid == 0) idx_out[cloud * s] = start;
    __syncthreads();
    for (int it = 1; it < s; ++it) {
        const float cx = xc[3 * cur], cy = xc[3 * cur + 1], cz = xc[3 * cur + 2];
        unsigned long long best = 0ull;
        for (int64_t i = tid; i < n; i += 1024) {
            const float dd = sqdist_nofma(xc[3 * i], xc[3 * i + 1], xc[3 * i + 2], cx, cy, cz);
            const float m = fminf(dc[i], dd);
            dc[i] = m;
            const unsigned long long key = ((unsigned long long)__float_as_uint(m) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)i);
            best = key > best ? key : best;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if (lane == 0) wkey[warp] = best;
        __syncthreads();
        if (warp == 0) {
            unsigned long long k = wkey[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o);
                k = other > k ? other : k;
            }
            if (lane == 0) bkey = k;
        }
        __syncthreads();
        cur = (int64_t)(0xFFFFFFFFu - (unsigned)bkey);
        if (tid == 0) idx_out[cloud * s + it] = (int32_t)cur;
        __syncthreads();
    }
}

template <int PPT, bool kProf = false, int SD = 0>
static int launch_fps(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, int cs, int64_t ppc, cudaStream_t st, int srows = 0)
{
    static bool attr_set = false;
    if (!attr_set) {
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_cluster_kernel<PPT, kProf, SD>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_cluster_kernel<PPT, kProf, SD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpsMaxSmemRows * kFpsThreads * 16));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((int64_t)b * cs));
    cfg.blockDim = dim3(kFpsThreads);
    cfg.dynamicSmemBytes = SD > 0 ? (size_t)((srows + 3) & ~3) * kFpsThreads * 12 : (size_t)srows * kFpsThreads * 16;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    static const int hw_barrier = getenv("PCP_FPS_HW_BARRIER") != nullptr;          // the previous exchange, kept for comparison
    PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, fps_cluster_kernel<PPT, kProf, SD>, d_x, n, s, start, d_idx, cs, ppc, hw_barrier, srows));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

static float *g_fps_scratch = nullptr;
static size_t g_fps_scratch_bytes = 0;

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_fps_f32(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_idx && b >= 0 && n >= 1 && s >= 1 && start >= 0 && start < n, PCP_ERR_INVALID,
                "pcp_fps_f32: bad arguments (b=%d n=%lld s=%d start=%d)", b, (long long)n, s, start);
    PCP_REQUIRE(n < 0xFFFFFFFFll, PCP_ERR_UNSUPPORTED, "pcp_fps_f32: n too large");
    if (b == 0) return PCP_OK;
    cudaStream_t st = as_stream(stream);
    // smallest cluster whose register file holds the cloud: cs * 512 threads * PPT points, PPT <= 25.
    // Prefer the widest PPT that fits (fewest CTAs per cloud => most clouds in flight), but widen the
    // cluster while the grid would leave SMs idle.
    constexpr int kMaxPPT = 25;
    int cs = 1;
    while (cs < kFpsMaxCluster && (n > (int64_t)cs * kFpsThreads * kMaxPPT)) cs *= 2;
    while (cs < 8 && (int64_t)b * cs * 2 <= sm_count() && n / cs > 2 * kFpsThreads) cs *= 2;
    if (n > (int64_t)cs * kFpsThreads * kMaxPPT) {
        const size_t need = (size_t)b * (size_t)n * sizeof(float);
        if (need > g_fps_scratch_bytes) {
            if (g_fps_scratch) cudaFree(g_fps_scratch);
            g_fps_scratch = nullptr; g_fps_scratch_bytes = 0;
            PCP_CUDA_OK(cudaMalloc(&g_fps_scratch, need));
            g_fps_scratch_bytes = need;
        }
        fps_stream_kernel<<<b, 1024, 0, st>>>(d_x, n, s, start, d_idx, g_fps_scratch);
        PCP_CHECK_LAUNCH();
        return PCP_OK;
    }
    // More clusters than the GPU holds at once (64 clouds x 8 CTAs on 148 SMs = four waves): halve the cluster and keep the
    // second half of every CTA's slice in shared memory (SoA, 16 B per point) beside the 25 rows in registers — half the
    // waves for a longer iteration.
    static const bool no_hybrid = getenv("PCP_FPS_NO_HYBRID") != nullptr;
    if (!no_hybrid && cs >= 2 && (int64_t)b * cs > sm_count()) {
        const int csh = cs / 2;
        const int64_t ppch = (n + csh - 1) / csh;
        const int rows = (int)((ppch + kFpsThreads - 1) / kFpsThreads);
        // 17 rows with everything in registers + up to 32 rows whose coordinates sit in shared memory and whose distances stay in
        // registers (17 * 4 + 32 = 100 data registers, 32 * 512 * 12 B = 192 KB)
        static const bool no_regd = getenv("PCP_FPS_NO_REGD") != nullptr;
        if (!no_regd && rows > kFpsRegRows && rows - kFpsRegRows <= kFpsRegDist) {
            static const bool prof_r = getenv("PCP_FPS_PROF") != nullptr;
            if (prof_r) return launch_fps<kFpsRegRows, true, kFpsRegDist>(d_x, b, n, s, start, d_idx, csh, ppch, st, rows - kFpsRegRows);
            return launch_fps<kFpsRegRows, false, kFpsRegDist>(d_x, b, n, s, start, d_idx, csh, ppch, st, rows - kFpsRegRows);
        }
        if (rows > kMaxPPT && rows - kMaxPPT <= kFpsMaxSmemRows) {
            static const bool prof_h = getenv("PCP_FPS_PROF") != nullptr;
            if (prof_h) return launch_fps<25, true>(d_x, b, n, s, start, d_idx, csh, ppch, st, rows - kMaxPPT);
            return launch_fps<25>(d_x, b, n, s, start, d_idx, csh, ppch, st, rows - kMaxPPT);
        }
    }
    const int64_t ppc = (n + cs - 1) / cs;
    const int ppt = (int)((ppc + kFpsThreads - 1) / kFpsThreads);
    if (ppt <= 1) return launch_fps<1>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    if (ppt <= 2) return launch_fps<2>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    if (ppt <= 4) return launch_fps<4>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    if (ppt <= 8) return launch_fps<8>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    if (ppt <= 13) return launch_fps<13>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    if (ppt <= 16) return launch_fps<16>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    if (ppt <= 20) return launch_fps<20>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    static const bool prof = getenv("PCP_FPS_PROF") != nullptr;
    if (prof) return launch_fps<25, true>(d_x, b, n, s, start, d_idx, cs, ppc, st);
    return launch_fps<25>(d_x, b, n, s, start, d_idx, cs, ppc, st);
}

/* development aid: cycles block 0 / thread 0 of the last PCP_FPS_PROF=1 launch (PPT = 25) spent in: distance loop, candidate
 * search + record, __syncthreads, warp-0 stage + DSMEM pushes, cluster barrier, slot read-back */
extern "C" int pcp_fps_phase_cycles(unsigned long long out[8])
{
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpyFromSymbol(out, g_fps_phase, 8 * sizeof(unsigned long long)));
    return PCP_OK;
}
