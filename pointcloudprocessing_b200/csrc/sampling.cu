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
//
// Round 2 adds the PRUNED variant behind pcp_fps_ws_f32 (caller-owned workspace): the same samples with ~90 % of the distance
// updates skipped — Morton-ordered cloud, per-group bounding spheres, warp-uniform skip; see the comment above fps_lean_kernel.
#include <cooperative_groups.h>

#include <algorithm>
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

// ------------------------------------------------------------------------------------------------------------------------------
// PRUNED sampler (round 2; fps_lean_kernel below).  The distance update of an iteration only changes points closer to the new
// sample than their running minimum; late in the chain that is a small neighbourhood.  The cloud is first put in Morton order of a
// coarse grid (a counting sort, fps_sort_* below, which only produces the permutation) and cut into groups of kFpsGroupRows rows of 32
// = 224 consecutive sorted points, dealt round-robin over the blocks of the cluster and their warps.  Per group: a bounding sphere
// and the group's largest running distance; a group whose sphere is farther from the new sample than sqrt(that distance) cannot
// change and is skipped by a warp-uniform branch.  Skipping is conservative (1e-4 relative slack against fp32 rounding of the test
// itself) and an updated group runs exactly the arithmetic of the unpruned kernel, so the samples stay bit-identical with the oracle;
// keys inside the kernel carry a block-local id and an exact "several points attain this maximum" flag, and only when the
// cluster-wide maximum is shared does an iteration fall back to the lowest ORIGINAL index through the permutation.
// (History: a first version kept the coordinates on chip — 13 rows in registers, 36 in shared memory, clusters of four, two waves
//  for 64 clouds: 3.26 ms; measurements of every step in profiles/r02_fps_pruned.txt.)
constexpr int kFpsGroupRows = 7;                      // rows of 32 points per group
constexpr int kFpsSortBits = 5;                       // grid of 32^3 cells over the cloud's bounding cube
constexpr int kFpsSortBins = 1 << (3 * kFpsSortBits);

__device__ __forceinline__ unsigned fps_fenc(float f) { const unsigned b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float fps_fdec(unsigned k) { return __uint_as_float((k & 0x80000000u) ? (k ^ 0x80000000u) : ~k); }
__device__ __forceinline__ unsigned fps_part1by2(unsigned v)
{
    v &= 0x3ffu; v = (v | (v << 16)) & 0x30000ffu; v = (v | (v << 8)) & 0x300f00fu; v = (v | (v << 4)) & 0x30c30c3u; v = (v | (v << 2)) & 0x9249249u;
    return v;
}
__device__ __forceinline__ unsigned fps_cell_of(float px, float py, float pz, float lx, float ly, float lz, float inv)
{   // NaN / Inf coordinates clamp into the grid (fmaxf drops a NaN): any cell is correct, the order only steers the pruning
    const float top = (float)((1 << kFpsSortBits) - 1);
    const unsigned ix = (unsigned)fminf(fmaxf((px - lx) * inv, 0.f), top), iy = (unsigned)fminf(fmaxf((py - ly) * inv, 0.f), top),
                   iz = (unsigned)fminf(fmaxf((pz - lz) * inv, 0.f), top);
    return fps_part1by2(ix) | (fps_part1by2(iy) << 1) | (fps_part1by2(iz) << 2);
}

__global__ void fps_sort_init_kernel(unsigned *__restrict__ bbox, int b, int *__restrict__ hist, int64_t nh)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (int64_t)b * 6) bbox[i] = (i % 6) < 3 ? 0xFFFFFFFFu : 0u;
    for (int64_t j = i; j < nh; j += (int64_t)gridDim.x * blockDim.x) hist[j] = 0;
}

// All global atomics of the sort go to a handful of words per cloud; issued per warp or per point they serialise in the L2 (bounding
// box: 447 k atomics on 384 words = 106 us; histogram / scatter: 7 M atomics on a very uneven cell population = 158 + 196 us).
// So: the bounding box is reduced per BLOCK first, and the histogram of a block's part of the cloud is built in SHARED memory
// (all 32 768 cells: 128 KB) — only per-bin totals / base claims go to global memory.
__global__ void __launch_bounds__(1024)
fps_sort_bbox_kernel(const float *__restrict__ x, int64_t n, unsigned *__restrict__ bbox)
{
    __shared__ unsigned slo[3][32], shi[3][32];
    const int cloud = blockIdx.y;
    const float *xc = x + (int64_t)cloud * n * 3;
    const int64_t nf = 3 * n;
    // flat, coalesced float loads; element e of a trip sits at i + 1024 e, its axis is (i + 1024 e) % 3 = (a0 + e) % 3
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};       // by (axis - a0) mod 3
    for (int64_t i = (int64_t)blockIdx.x * 3072 + tid; i < nf; i += (int64_t)gridDim.x * 3072) {
#pragma unroll
        for (int e = 0; e < 3; ++e) {
            const int64_t j = i + 1024 * e;
            if (j < nf) { const unsigned k = fps_fenc(__ldg(xc + j)); lo[e] = min(lo[e], k); hi[e] = max(hi[e], k); }
        }
    }
    // (the stride 3072 * gridDim.x is a multiple of 3, so a thread's slot e always holds axis (tid + 1024 e + 3072 blockIdx.x) % 3
    //  = (tid + e) % 3)
    const int a0 = tid % 3;
    unsigned alo[3], ahi[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const int e = (a - a0 + 3) % 3;
        alo[a] = e == 0 ? lo[0] : e == 1 ? lo[1] : lo[2];
        ahi[a] = e == 0 ? hi[0] : e == 1 ? hi[1] : hi[2];
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const unsigned l = __reduce_min_sync(0xffffffffu, alo[a]), h = __reduce_max_sync(0xffffffffu, ahi[a]);
        if (lane == 0) { slo[a][warp] = l; shi[a][warp] = h; }
    }
    __syncthreads();
    if (warp < 3) {
        const unsigned l = __reduce_min_sync(0xffffffffu, slo[warp][lane]), h = __reduce_max_sync(0xffffffffu, shi[warp][lane]);
        if (lane == 0) { atomicMin(&bbox[cloud * 6 + warp], l); atomicMax(&bbox[cloud * 6 + 3 + warp], h); }
    }
}

// blockIdx.x = part of the cloud (gridDim.x parts of `per` points), blockIdx.y = cloud; dynamic shared memory: kFpsSortBins ints.
// kScatter = false: hist[cell] += this part's count.  kScatter = true (hist holds the exclusive scan): claim a base per cell for
// this part (one global atomic per occupied cell), then place the part's points behind it with shared-memory cursors.
template <bool kScatter>
__global__ void __launch_bounds__(1024)
fps_sort_pass_kernel(const float *__restrict__ x, int64_t n, int64_t per, const unsigned *__restrict__ bbox, int *__restrict__ hist,
                     int32_t *__restrict__ perm)
{
    extern __shared__ int fps_cnt[];
    const int cloud = blockIdx.y, tid = threadIdx.x;
    const float *xc = x + (int64_t)cloud * n * 3;
    int *h = hist + (int64_t)cloud * kFpsSortBins;
    const int64_t p0 = (int64_t)blockIdx.x * per, p1 = min(n, p0 + per);
    const unsigned *bb = bbox + cloud * 6;
    const float lx = fps_fdec(bb[0]), ly = fps_fdec(bb[1]), lz = fps_fdec(bb[2]);
    const float ext = fmaxf(fmaxf(fps_fdec(bb[3]) - lx, fps_fdec(bb[4]) - ly), fps_fdec(bb[5]) - lz);
    const float inv = ext > 0.f ? (float)(1 << kFpsSortBits) / ext : 0.f;
    for (int i = tid; i < kFpsSortBins; i += 1024) fps_cnt[i] = 0;
    __syncthreads();
    for (int64_t i0 = p0 + tid; i0 < p1; i0 += 4096) {
        unsigned c[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int64_t i = i0 + 1024 * e;
            c[e] = i < p1 ? fps_cell_of(__ldg(xc + 3 * i), __ldg(xc + 3 * i + 1), __ldg(xc + 3 * i + 2), lx, ly, lz, inv) : 0xFFFFFFFFu;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (c[e] != 0xFFFFFFFFu) atomicAdd(&fps_cnt[c[e]], 1);
    }
    __syncthreads();
    if constexpr (!kScatter) {
        for (int i = tid; i < kFpsSortBins; i += 1024) { const int v = fps_cnt[i]; if (v) atomicAdd(&h[i], v); }
    } else {
        for (int i = tid; i < kFpsSortBins; i += 1024) { const int v = fps_cnt[i]; fps_cnt[i] = v ? atomicAdd(&h[i], v) : 0; }
        __syncthreads();
        int32_t *pc = perm + (int64_t)cloud * n;
        for (int64_t i0 = p0 + tid; i0 < p1; i0 += 4096) {
            unsigned c[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int64_t i = i0 + 1024 * e;
                c[e] = i < p1 ? fps_cell_of(__ldg(xc + 3 * i), __ldg(xc + 3 * i + 1), __ldg(xc + 3 * i + 2), lx, ly, lz, inv) : 0xFFFFFFFFu;
            }
            // only the permutation is written: a sorted COPY would be 12-byte writes at random positions (partial sectors); the
            // sampler gathers its points through the permutation instead
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (c[e] != 0xFFFFFFFFu) pc[atomicAdd(&fps_cnt[c[e]], 1)] = (int32_t)(i0 + 1024 * e);
        }
    }
}

__global__ void __launch_bounds__(1024)
fps_sort_scan_kernel(int *__restrict__ hist)
{   // exclusive scan of one cloud's kFpsSortBins counters, in place; 1024 bins per round, coalesced
    __shared__ int wsum[32];
    __shared__ int carry_s;
    int *h = hist + (int64_t)blockIdx.x * kFpsSortBins;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int carry = 0;
    for (int base = 0; base < kFpsSortBins; base += 1024) {
        const int v = h[base + tid];
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const int w = wsum[lane];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
            wsum[lane] = wi - w;
            if (lane == 31) carry_s = wi;
        }
        __syncthreads();
        h[base + tid] = carry + wsum[warp] + inc - v;
        carry += carry_s;
        __syncthreads();
    }
}

__device__ __forceinline__ float fps_skip_threshold(float rad_inflated, float dmax)
{   // skip when |q - c|^2 (fp32, any rounding) > thr: then every point of the sphere is farther from q than sqrt(dmax).
    // sqrt.approx is within a few ulp — far inside the 1e-4 slack
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(dmax));
    const float t = rad_inflated + r * 1.0001f;
    return t * t * 1.0002f;
}

// the arg-max of the keys held by the lanes (low word 0: no key): its high word, the lane that holds the first of them, and whether it
// is shared
__device__ __forceinline__ unsigned fps_key_max(unsigned hi, unsigned lo, int &sel, bool &multi)
{
    const unsigned mhi = __reduce_max_sync(0xffffffffu, lo ? hi : 0u);
    const unsigned eq = __ballot_sync(0xffffffffu, lo != 0u && hi == mhi);
    sel = eq ? __ffs(eq) - 1 : -1;
    const unsigned slo = __shfl_sync(0xffffffffu, lo, sel < 0 ? 0 : sel);
    multi = __popc(eq) > 1 || (sel >= 0 && (slo & 0x40000000u));
    return mhi;
}

// Exchange of the pruned sampler: there is no block stage.  Every WARP pushes ONE 16-byte record (x, y, z, code) into every block of
// the cluster (st.async onto the receiver's transaction barrier) and every warp reduces the cs * 16 records itself; warps of a block
// never meet.  code = 0: the warp holds no point; else (float bits of its largest running distance + 1) | bit 31 when several of its
// points attain it.  The per-group state (sphere, threshold, key, coordinates of the point behind the key) lives in REGISTERS of
// lane g: the skip tests of all groups of a warp are one compare and a ballot.
constexpr int kFpsPruneMaxCluster = 8;
// try_wait with a suspend-time hint: the waiter sleeps in hardware instead of coming back every few hundred cycles to spin
__device__ __forceinline__ void fps_wait_records(uint64_t *bar, uint32_t parity)
{
    uint32_t ok = 0u;
    const long long t0 = clock64();
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(umma::smem_u32(bar)), "r"(parity), "r"(20000u) : "memory");
        if (ok) return;
        if (clock64() - t0 > 4000000000ll) __trap();       // a lost record must fail the launch, not hang the device
    }
}

// ------------------------------------------------------------------------------------------------------------------------------
// LEAN pruned variant: only the running DISTANCES stay on chip.  With pruning ~9 % of the groups need their coordinates in an
// iteration, so the coordinates can live in an L2-resident copy (per thread row-major in 16-byte chunks [group][chunk][tid]: a group
// is six coalesced LDG.128) and a block only has to hold 4 bytes per point — 98 rows of 512 distances in shared memory = 50 176
// points per block.  A 100 000-point cloud then fits in a cluster of TWO, 64 clouds are 128 blocks = ONE wave on 148 SMs instead of
// the two waves of the register / shared-memory resident kernels, and the whole kernel is generic code (no register-resident rows
// to unroll over): the loop body is a few KB.
constexpr int kLeanMaxGroups = 14;                    // 14 x 7 rows x 512 x 4 B = 200 704 B of distances
constexpr int kLeanChunks = 6;                        // 7 rows x 3 floats = 21 (+ 3 padding) floats = 6 float4 per thread and group

__device__ __forceinline__ unsigned fps_lean_id(bool multi, int warp, int g, int j, int lane)
{
    return 0x80000000u | (multi ? 0x40000000u : 0u) | ((unsigned)warp << 12) | ((unsigned)g << 8) | ((unsigned)j << 5) | (unsigned)lane;
}

template <int CS, bool kProf>
__global__ void __launch_bounds__(kFpsThreads, 1)
fps_lean_kernel(const int32_t *__restrict__ perm, const float *__restrict__ xorig, float4 *__restrict__ coords, int64_t n, int s, int start,
                int32_t *__restrict__ idx_out, int ngroups, int dbg)
{
    constexpr int cs = CS;
    constexpr int nrec = CS * kFpsWarps;
    constexpr int kSlots = nrec < 32 ? 32 : nrec;
    constexpr int kPerLane = nrec <= 32 ? 1 : nrec / 32;
    extern __shared__ __align__(16) float fps_dsm[];              // running distances [ngroups * 7][512]
    __shared__ float4 slots[2][kSlots], tslots[2][kSlots];
    __shared__ uint64_t xbar[2], tbar[2];

    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int64_t cloud = blockIdx.x / cs;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float *xo = xorig + cloud * n * 3;
    const int32_t *pc = perm + cloud * n;
    float4 *cb = coords + (size_t)blockIdx.x * ngroups * kLeanChunks * kFpsThreads + tid;      // this thread's chunk c of group g: cb[(g * 6 + c) * 512]
    // group slot g of this warp holds sorted group ((g * 16 + warp) * cs + rank): 224 consecutive sorted points, dealt round-robin
    auto gpos = [&](int g, int j) -> int64_t { return ((int64_t)((g * kFpsWarps + warp) * cs + rank) * kFpsGroupRows + j) * 32 + lane; };
    auto real_rows = [&](int g) -> unsigned {      // bit j: row j of group g is a point of the cloud (rows fill up in order)
        const int64_t left = n - gpos(g, 0);
        const int cnt = left <= 0 ? 0 : (int)min((int64_t)kFpsGroupRows, (left + 31) / 32);
        return (1u << cnt) - 1u;
    };

    // ---- set-up: gather the points through the permutation, write the coordinate copy, clear the distances, bounding spheres
    float scx = 0.f, scy = 0.f, scz = 0.f, srad = 0.f, sthr = -1.f;      // lane g: sphere of group g; thr -1: always skipped
    float gx = 0.f, gy = 0.f, gz = 0.f;                                   // lane g: coordinates of the point behind the group's key
    unsigned khi = 0u, klo = 0u;                                          // lane g: key of group g (klo: fps_lean_id, 0 = no point)
    for (int g = 0; g < ngroups; ++g) {
        const unsigned mine = real_rows(g);
        float f[24];
#pragma unroll
        for (int j = 0; j < kFpsGroupRows; ++j) {
            const bool real = (mine >> j) & 1u;
            const int64_t o = real ? (int64_t)__ldg(pc + gpos(g, j)) : 0;
            f[3 * j] = real ? __ldg(xo + 3 * o) : 0.f; f[3 * j + 1] = real ? __ldg(xo + 3 * o + 1) : 0.f; f[3 * j + 2] = real ? __ldg(xo + 3 * o + 2) : 0.f;
            fps_dsm[(g * kFpsGroupRows + j) * kFpsThreads + tid] = real ? __int_as_float(0x7f800000) : 0.f;
        }
        f[21] = f[22] = f[23] = 0.f;
#pragma unroll
        for (int c = 0; c < kLeanChunks; ++c) cb[(size_t)(g * kLeanChunks + c) * kFpsThreads] = make_float4(f[4 * c], f[4 * c + 1], f[4 * c + 2], f[4 * c + 3]);
        float ax = 0.f, ay = 0.f, az = 0.f, cnt = 0.f;
#pragma unroll
        for (int j = 0; j < kFpsGroupRows; ++j)
            if ((mine >> j) & 1u) { ax += f[3 * j]; ay += f[3 * j + 1]; az += f[3 * j + 2]; cnt += 1.f; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ax += __shfl_xor_sync(0xffffffffu, ax, o); ay += __shfl_xor_sync(0xffffffffu, ay, o);
            az += __shfl_xor_sync(0xffffffffu, az, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        }
        const float ic = cnt > 0.f ? 1.f / cnt : 0.f;
        const float mx = ax * ic, my = ay * ic, mz = az * ic;
        float r2 = 0.f;
#pragma unroll
        for (int j = 0; j < kFpsGroupRows; ++j)
            if ((mine >> j) & 1u)
                r2 = fmaxf(r2, (f[3 * j] - mx) * (f[3 * j] - mx) + (f[3 * j + 1] - my) * (f[3 * j + 1] - my) + (f[3 * j + 2] - mz) * (f[3 * j + 2] - mz));
        r2 = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(r2)));
        if (lane == g) {
            scx = mx; scy = my; scz = mz;
            srad = sqrtf(r2) * 1.0001f + 1e-30f;
            sthr = cnt > 0.f ? __int_as_float(0x7f800000) : -1.f;      // a populated group is never skipped before its first update
        }
    }
    float cx = __ldg(xo + 3 * (int64_t)start), cy = __ldg(xo + 3 * (int64_t)start + 1), cz = __ldg(xo + 3 * (int64_t)start + 2);
    if (rank == 0 && tid == 0) idx_out[cloud * s] = start;
    if (tid < kSlots) {
        slots[0][tid] = make_float4(0.f, 0.f, 0.f, 0.f); slots[1][tid] = slots[0][tid];
        tslots[0][tid] = make_float4(0.f, 0.f, 0.f, __uint_as_float(0xFFFFFFFFu)); tslots[1][tid] = tslots[0][tid];
    }
    if (tid == 0) {
        umma::mbar_init(&xbar[0], 1); umma::mbar_init(&xbar[1], 1); umma::mbar_init(&tbar[0], 1); umma::mbar_init(&tbar[1], 1);
        umma::fence_barrier_init();
    }
    __syncthreads();       // (also: this thread's coordinate copy is only ever read back by itself)
    if (tid == 0) {
        umma::mbar_arrive_expect_tx(&xbar[0], nrec * 16u); umma::mbar_arrive_expect_tx(&xbar[1], nrec * 16u);
        umma::mbar_arrive_expect_tx(&tbar[0], nrec * 16u); umma::mbar_arrive_expect_tx(&tbar[1], nrec * 16u);
    }
    cluster.sync();                              // every block is armed before anybody pushes
    long long ph[6] = {0, 0, 0, 0, 0, 0}, tprev = clock64();
    unsigned long long nact = 0, nties = 0;
    unsigned ntie = 0u;                          // tie iterations so far (identical in every warp of the cluster)
    unsigned wcode = 0u, wlo = 0u;               // this warp's record (warp-uniform)
    float wx = 0.f, wy = 0.f, wz = 0.f;
    auto load_group = [&](float4 (&c)[kLeanChunks], int g) {      // plain loads: the copy was written by this kernel
#pragma unroll
        for (int k = 0; k < kLeanChunks; ++k) c[k] = cb[(size_t)(g * kLeanChunks + k) * kFpsThreads];
    };
    for (int it = 1; it < s; ++it) {
        const float tdx = cx - scx, tdy = cy - scy, tdz = cz - scz;
        unsigned act = __ballot_sync(0xffffffffu, !(tdx * tdx + tdy * tdy + tdz * tdz > sthr));
        if (dbg == 1 && it > 1) act = 0u;            // development: exchange only (results wrong)
        if (act) {
            unsigned rem = act;
            int g = __ffs(rem) - 1;
            rem &= rem - 1u;
            float4 c[kLeanChunks];
            load_group(c, g);
            while (true) {
                // the next active group's coordinates are in flight while this one is updated
                const int gn = rem ? __ffs(rem) - 1 : -1;
                float4 cn[kLeanChunks];
                if (gn >= 0) { rem &= rem - 1u; load_group(cn, gn); }
                if (kProf) ++nact;
                const float f[24] = {c[0].x, c[0].y, c[0].z, c[0].w, c[1].x, c[1].y, c[1].z, c[1].w, c[2].x, c[2].y, c[2].z, c[2].w,
                                     c[3].x, c[3].y, c[3].z, c[3].w, c[4].x, c[4].y, c[4].z, c[4].w, c[5].x, c[5].y, c[5].z, c[5].w};
                float *dg = fps_dsm + (size_t)g * kFpsGroupRows * kFpsThreads + tid;
                float m[kFpsGroupRows], lbv = -1.f;
#pragma unroll
                for (int j = 0; j < kFpsGroupRows; ++j) {
                    m[j] = fminf(dg[j * kFpsThreads], sqdist_nofma(f[3 * j], f[3 * j + 1], f[3 * j + 2], cx, cy, cz));
                    dg[j * kFpsThreads] = m[j];
                    lbv = fmaxf(lbv, m[j]);
                }
                // the group's maximum over its REAL points (padding rows hold 0 and never raise it), the first point that attains it,
                // whether it is alone, and that point's coordinates (kept with the key: a later record needs no fetch)
                const unsigned mine = real_rows(g);
                const unsigned gm = __reduce_max_sync(0xffffffffu, mine ? __float_as_uint(lbv) : 0u);
                unsigned at = 0u;
#pragma unroll
                for (int j = 0; j < kFpsGroupRows; ++j)
                    if (__float_as_uint(m[j]) == gm) at |= 1u << j;
                at &= mine;
                const unsigned lanes_at = __ballot_sync(0xffffffffu, at != 0u);
                const unsigned multi_l = __ballot_sync(0xffffffffu, (at & (at - 1u)) != 0u);
                const int ol = lanes_at ? __ffs(lanes_at) - 1 : 0;
                const unsigned oat = __shfl_sync(0xffffffffu, at, ol);
                const int oj = oat ? __ffs(oat) - 1 : 0;
                float ox = 0.f, oy = 0.f, oz = 0.f;
#pragma unroll
                for (int j = 0; j < kFpsGroupRows; ++j)
                    if (j == oj) { ox = f[3 * j]; oy = f[3 * j + 1]; oz = f[3 * j + 2]; }
                ox = __shfl_sync(0xffffffffu, ox, ol); oy = __shfl_sync(0xffffffffu, oy, ol); oz = __shfl_sync(0xffffffffu, oz, ol);
                if (lane == g) {
                    khi = gm;
                    klo = lanes_at ? fps_lean_id((lanes_at & (lanes_at - 1u)) != 0u || multi_l != 0u, warp, g, oj, ol) : 0u;
                    sthr = lanes_at ? fps_skip_threshold(srad, __uint_as_float(gm)) : -1.f;
                    gx = ox; gy = oy; gz = oz;
                }
                if (gn < 0) break;
                g = gn;
#pragma unroll
                for (int k = 0; k < kLeanChunks; ++k) c[k] = cn[k];
            }
            // the warp's record: its best group's key and that point's coordinates
            int sel; bool multi;
            const unsigned whi = fps_key_max(khi, klo, sel, multi);
            wlo = sel < 0 ? 0u : __shfl_sync(0xffffffffu, klo, sel);
            wcode = sel < 0 ? 0u : (whi + 1u) | (multi ? 0x80000000u : 0u);
            const int ss = sel < 0 ? 0 : sel;
            wx = __shfl_sync(0xffffffffu, gx, ss); wy = __shfl_sync(0xffffffffu, gy, ss); wz = __shfl_sync(0xffffffffu, gz, ss);
        }
        FPS_STAMP(0);
        const int par = it & 1;
        if (lane < cs)       // lane t -> block t of the cluster: data and its arrival in one remote operation
            fps_st_async16(&slots[par][rank * kFpsWarps + warp], &xbar[par], (uint32_t)lane, __float_as_uint(wx), __float_as_uint(wy),
                           __float_as_uint(wz), wcode);
        FPS_STAMP(3);
        fps_wait_records(&xbar[par], ((unsigned)(it - 1) >> 1) & 1u);
        if (tid == 0 && it + 2 < s) umma::mbar_arrive_expect_tx(&xbar[par], nrec * 16u);      // re-arm for iteration it + 2
        FPS_STAMP(4);
        // every warp reduces the records itself
        float4 rc[kPerLane];
#pragma unroll
        for (int e = 0; e < kPerLane; ++e) rc[e] = slots[par][lane + 32 * e];
        float4 brec = rc[0];
        unsigned bcode = __float_as_uint(rc[0].w), bi = (unsigned)lane;
#pragma unroll
        for (int e = 1; e < kPerLane; ++e) {
            const unsigned cc = __float_as_uint(rc[e].w);
            if ((cc & 0x7FFFFFFFu) > (bcode & 0x7FFFFFFFu)) { brec = rc[e]; bcode = cc; bi = (unsigned)(lane + 32 * e); }
            else if (cc != 0u && (cc & 0x7FFFFFFFu) == (bcode & 0x7FFFFFFFu)) bcode |= 0x80000000u;
        }
        const unsigned cmax = __reduce_max_sync(0xffffffffu, bcode & 0x7FFFFFFFu);       // > 0: the cloud has points
        const unsigned eq = __ballot_sync(0xffffffffu, (bcode & 0x7FFFFFFFu) == cmax && bcode != 0u);
        const int sl = __ffs(eq) - 1;
        const bool tie = (eq & (eq - 1u)) != 0u || (__shfl_sync(0xffffffffu, bcode, sl) & 0x80000000u) != 0u;
        const unsigned dmax = cmax - 1u;
        if (!tie) {
            cx = __shfl_sync(0xffffffffu, brec.x, sl); cy = __shfl_sync(0xffffffffu, brec.y, sl); cz = __shfl_sync(0xffffffffu, brec.z, sl);
            const int wi = (int)__shfl_sync(0xffffffffu, bi, sl);
            // the one thread that holds the winner translates its sorted position (a load that nothing in the loop waits for)
            if (wi == rank * kFpsWarps + warp && lane == (int)(wlo & 31u))
                idx_out[cloud * s + it] = __ldg(pc + gpos((int)((wlo >> 8) & 15u), (int)((wlo >> 5) & 7u)));
        } else {
            // several points share the largest distance: np.argmax takes the lowest original index
            if (kProf) ++nties;
            unsigned cand = 0xFFFFFFFFu;
            int cg_ = 0, cj = 0;
            unsigned gat = __ballot_sync(0xffffffffu, klo != 0u && khi == dmax);      // groups of this warp at the maximum
            while (gat) {
                const int g = __ffs(gat) - 1;
                gat &= gat - 1u;
                const unsigned mine = real_rows(g);
                for (int j = 0; j < kFpsGroupRows; ++j)
                    if (((mine >> j) & 1u) && __float_as_uint(fps_dsm[(size_t)(g * kFpsGroupRows + j) * kFpsThreads + tid]) == dmax) {
                        const unsigned o = (unsigned)__ldg(pc + gpos(g, j));
                        if (o < cand) { cand = o; cg_ = g; cj = j; }
                    }
            }
            const unsigned wmin = __reduce_min_sync(0xffffffffu, cand);
            float rx = 0.f, ry = 0.f, rz = 0.f;
            const unsigned who = __ballot_sync(0xffffffffu, wmin != 0xFFFFFFFFu && cand == wmin);
            const int ol = who ? __ffs(who) - 1 : 0;
            if (who && lane == ol) {
                const float *fb = reinterpret_cast<const float *>(cb);
                auto comp = [&](int q) { return fb[((size_t)(cg_ * kLeanChunks + (q >> 2)) * kFpsThreads) * 4 + (q & 3)]; };
                rx = comp(3 * cj); ry = comp(3 * cj + 1); rz = comp(3 * cj + 2);
            }
            rx = __shfl_sync(0xffffffffu, rx, ol); ry = __shfl_sync(0xffffffffu, ry, ol); rz = __shfl_sync(0xffffffffu, rz, ol);
            const int tpar = (int)(ntie & 1u);
            if (lane < cs)
                fps_st_async16(&tslots[tpar][rank * kFpsWarps + warp], &tbar[tpar], (uint32_t)lane, __float_as_uint(rx), __float_as_uint(ry),
                               __float_as_uint(rz), wmin);
            fps_wait_records(&tbar[tpar], (ntie >> 1) & 1u);
            if (tid == 0) umma::mbar_arrive_expect_tx(&tbar[tpar], nrec * 16u);
            ++ntie;
            float4 tr[kPerLane];
#pragma unroll
            for (int e = 0; e < kPerLane; ++e) tr[e] = tslots[tpar][lane + 32 * e];
            float4 tb4 = tr[0];
#pragma unroll
            for (int e = 1; e < kPerLane; ++e)
                if (__float_as_uint(tr[e].w) < __float_as_uint(tb4.w)) tb4 = tr[e];
            const unsigned tmin = __reduce_min_sync(0xffffffffu, __float_as_uint(tb4.w));
            const int tb = __ffs(__ballot_sync(0xffffffffu, __float_as_uint(tb4.w) == tmin)) - 1;
            cx = __shfl_sync(0xffffffffu, tb4.x, tb); cy = __shfl_sync(0xffffffffu, tb4.y, tb); cz = __shfl_sync(0xffffffffu, tb4.z, tb);
            if (rank == 0 && tid == 0) idx_out[cloud * s + it] = (int32_t)tmin;
        }
        FPS_STAMP(5);
    }
    if (kProf) {
        if (blockIdx.x == 0 && lane == 0) atomicAdd(&g_fps_phase[6], nact);
        if (blockIdx.x == 0 && tid == 0) {
            for (int i = 0; i < 6; ++i) g_fps_phase[i] = (unsigned long long)ph[i];
            g_fps_phase[7] = nties;
        }
    }
    cluster.sync();     // no block may exit while a peer can still write into its slots
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

static inline int fps_device()
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return -1;
    return dev;
}

template <int PPT, bool kProf = false, int SD = 0>
static int launch_fps(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, int cs, int64_t ppc, cudaStream_t st, int srows = 0)
{
    static bool attr_set[64] = {};          // per device (per template instance): function attributes belong to a context
    const int dev = fps_device();
    PCP_REQUIRE(dev >= 0, PCP_ERR_CUDA, "fps: no current device");
    if (!attr_set[dev]) {
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_cluster_kernel<PPT, kProf, SD>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_cluster_kernel<PPT, kProf, SD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpsMaxSmemRows * kFpsThreads * 16));
        attr_set[dev] = true;
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

static float *g_fps_scratch[64] = {};        // per device: running distances of the streaming fallback when the caller passes no workspace
static size_t g_fps_scratch_bytes[64] = {};
// clouds too large for the widest register-resident cluster take the streaming kernel (one float of scratch per point)
static inline bool fps_needs_stream(int64_t n) { return n > (int64_t)kFpsMaxCluster * kFpsThreads * 25; }
static int fps_plain(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, float *dist_ws, cudaStream_t st);

// ---- pruned path: geometry, workspace layout and launch
struct FpsPrunePlan {
    int cs = 0, ngroups = 0;
    size_t off_perm = 0, off_hist = 0, off_bbox = 0, off_coords = 0, bytes = 0;
};

static bool fps_prune_plan(int b, int64_t n, int s, FpsPrunePlan &pl)
{
    static const bool off = getenv("PCP_FPS_NO_PRUNE") != nullptr;
    if (off || s < 32 || n >= 0x7FFFFFFFll) return false;
    // the cluster whose blocks hold the cloud's running distances (50 176 points per block)
    int cs = 1;
    while (cs <= kFpsPruneMaxCluster && n > (int64_t)cs * kLeanMaxGroups * kFpsWarps * kFpsGroupRows * 32) cs *= 2;
    if (cs > kFpsPruneMaxCluster) return false;
    // only where the register-resident clusters (25 rows per block) would not all be resident at once — the batch sizes the
    // hybrid kernel was built for; a few clouds are faster spread over wide clusters
    int cs_reg = 1;
    while (cs_reg < kFpsMaxCluster && n > (int64_t)cs_reg * kFpsThreads * 25) cs_reg *= 2;
    static const bool force = getenv("PCP_FPS_FORCE_PRUNE") != nullptr;
    if (!force && (int64_t)b * cs_reg <= sm_count()) return false;
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const int64_t per_group = (int64_t)cs * kFpsWarps * kFpsGroupRows * 32;      // points one group slot holds across the cluster
    const int ngroups = (int)((n + per_group - 1) / per_group);
    pl.cs = cs; pl.ngroups = ngroups;
    pl.off_perm = 0;
    pl.off_hist = pl.off_perm + up((size_t)b * n * sizeof(int32_t));
    pl.off_bbox = pl.off_hist + up((size_t)b * kFpsSortBins * sizeof(int));
    pl.off_coords = pl.off_bbox + up((size_t)b * 6 * sizeof(unsigned));
    pl.bytes = pl.off_coords + up((size_t)b * cs * ngroups * kLeanChunks * kFpsThreads * sizeof(float4));
    return true;
}

template <int CS, bool kProf>
static int launch_fps_lean_cs(const int32_t *perm, const float *d_x, float4 *coords, int b, int64_t n, int s, int start, int32_t *d_idx,
                              const FpsPrunePlan &pl, int dbg, cudaStream_t st)
{
    auto kern = fps_lean_kernel<CS, kProf>;
    static bool attr_set_dev[64] = {};      // per device (per template instance)
    const int dev = fps_device();
    PCP_REQUIRE(dev >= 0, PCP_ERR_CUDA, "fps: no current device");
    if (!attr_set_dev[dev]) {
        PCP_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        PCP_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kLeanMaxGroups * kFpsGroupRows * kFpsThreads * 4));
        attr_set_dev[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((int64_t)b * CS));
    cfg.blockDim = dim3(kFpsThreads);
    cfg.dynamicSmemBytes = (size_t)pl.ngroups * kFpsGroupRows * kFpsThreads * 4;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, perm, d_x, coords, n, s, start, d_idx, pl.ngroups, dbg));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

template <bool kProf>
static int launch_fps_pruned(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, const FpsPrunePlan &pl, void *ws, cudaStream_t st)
{
    auto *base = static_cast<unsigned char *>(ws);
    int32_t *perm = reinterpret_cast<int32_t *>(base + pl.off_perm);
    int *hist = reinterpret_cast<int *>(base + pl.off_hist);
    unsigned *bbox = reinterpret_cast<unsigned *>(base + pl.off_bbox);
    const int64_t nh = (int64_t)b * kFpsSortBins;
    fps_sort_init_kernel<<<(unsigned)std::min<int64_t>((nh + 255) / 256, 4096), 256, 0, st>>>(bbox, b, hist, nh);
    PCP_CHECK_LAUNCH();
    // parts per cloud: one resident wave of blocks (a block holds the whole 128 KB histogram: one per SM)
    const int parts = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)16, (int64_t)sm_count() / b, (n + 8191) / 8192}));
    const int64_t per = (((n + parts - 1) / parts) + 3) & ~(int64_t)3;
    const dim3 grid((unsigned)parts, (unsigned)b);
    static bool attr_set[64] = {};
    const int sdev = fps_device();
    PCP_REQUIRE(sdev >= 0, PCP_ERR_CUDA, "fps: no current device");
    if (!attr_set[sdev]) {
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_sort_pass_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpsSortBins * 4));
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_sort_pass_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kFpsSortBins * 4));
        attr_set[sdev] = true;
    }
    const int bparts = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)32, (int64_t)4 * sm_count() / b, (3 * n + 3071) / 3072}));
    fps_sort_bbox_kernel<<<dim3((unsigned)bparts, (unsigned)b), 1024, 0, st>>>(d_x, n, bbox);
    PCP_CHECK_LAUNCH();
    fps_sort_pass_kernel<false><<<grid, 1024, kFpsSortBins * 4, st>>>(d_x, n, per, bbox, hist, nullptr);
    PCP_CHECK_LAUNCH();
    fps_sort_scan_kernel<<<b, 1024, 0, st>>>(hist);
    PCP_CHECK_LAUNCH();
    fps_sort_pass_kernel<true><<<grid, 1024, kFpsSortBins * 4, st>>>(d_x, n, per, bbox, hist, perm);
    PCP_CHECK_LAUNCH();

    static const int dbg = getenv("PCP_FPS_PRUNE_DBG") ? atoi(getenv("PCP_FPS_PRUNE_DBG")) : 0;      // development (see the kernel)
    if (dbg == 2) return PCP_OK;                     // development: the sort alone
    float4 *coords = reinterpret_cast<float4 *>(base + pl.off_coords);
    switch (pl.cs) {
    case 1: return launch_fps_lean_cs<1, kProf>(perm, d_x, coords, b, n, s, start, d_idx, pl, dbg, st);
    case 2: return launch_fps_lean_cs<2, kProf>(perm, d_x, coords, b, n, s, start, d_idx, pl, dbg, st);
    case 4: return launch_fps_lean_cs<4, kProf>(perm, d_x, coords, b, n, s, start, d_idx, pl, dbg, st);
    default: return launch_fps_lean_cs<8, kProf>(perm, d_x, coords, b, n, s, start, d_idx, pl, dbg, st);
    }
}

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_fps_f32(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, pcp_stream stream);

extern "C" int pcp_fps_workspace_bytes(int b, int64_t n, int s, size_t *bytes)
{
    PCP_REQUIRE(bytes && b >= 0 && n >= 1 && s >= 1, PCP_ERR_INVALID, "pcp_fps_workspace_bytes: bad arguments");
    FpsPrunePlan pl;
    if (b > 0 && fps_prune_plan(b, n, s, pl)) *bytes = pl.bytes;
    else if (b > 0 && fps_needs_stream(n)) *bytes = ((size_t)b * (size_t)n * sizeof(float) + 255) & ~(size_t)255;      // running distances
    else *bytes = 0;
    return PCP_OK;
}

extern "C" int pcp_fps_ws_f32(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, void *d_workspace, size_t workspace_bytes,
                              pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_idx && b >= 0 && n >= 1 && s >= 1 && start >= 0 && start < n, PCP_ERR_INVALID,
                "pcp_fps_ws_f32: bad arguments (b=%d n=%lld s=%d start=%d)", b, (long long)n, s, start);
    if (b == 0) return PCP_OK;
    FpsPrunePlan pl;
    PCP_REQUIRE(n < 0xFFFFFFFFll, PCP_ERR_UNSUPPORTED, "pcp_fps_ws_f32: n too large");
    if (!fps_prune_plan(b, n, s, pl)) {
        float *dist = nullptr;
        if (fps_needs_stream(n)) {
            PCP_REQUIRE(d_workspace && workspace_bytes >= (size_t)b * (size_t)n * sizeof(float), PCP_ERR_INVALID,
                        "pcp_fps_ws_f32: workspace of %zu bytes needed, %zu given — size it with pcp_fps_workspace_bytes",
                        (size_t)b * (size_t)n * sizeof(float), workspace_bytes);
            dist = static_cast<float *>(d_workspace);
        }
        return fps_plain(d_x, b, n, s, start, d_idx, dist, as_stream(stream));
    }
    PCP_REQUIRE(d_workspace && workspace_bytes >= pl.bytes && (reinterpret_cast<uintptr_t>(d_workspace) & 255) == 0, PCP_ERR_INVALID,
                "pcp_fps_ws_f32: workspace of %zu bytes (256-byte aligned) needed, %zu given — size it with pcp_fps_workspace_bytes", pl.bytes,
                workspace_bytes);
    static const bool prof = getenv("PCP_FPS_PROF") != nullptr;
    if (prof) {
        unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        PCP_CUDA_OK(cudaMemcpyToSymbol(g_fps_phase, zero, sizeof(zero)));
        return launch_fps_pruned<true>(d_x, b, n, s, start, d_idx, pl, d_workspace, as_stream(stream));
    }
    return launch_fps_pruned<false>(d_x, b, n, s, start, d_idx, pl, d_workspace, as_stream(stream));
}

extern "C" int pcp_fps_f32(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_idx && b >= 0 && n >= 1 && s >= 1 && start >= 0 && start < n, PCP_ERR_INVALID,
                "pcp_fps_f32: bad arguments (b=%d n=%lld s=%d start=%d)", b, (long long)n, s, start);
    PCP_REQUIRE(n < 0xFFFFFFFFll, PCP_ERR_UNSUPPORTED, "pcp_fps_f32: n too large");
    if (b == 0) return PCP_OK;
    return fps_plain(d_x, b, n, s, start, d_idx, nullptr, as_stream(stream));
}

namespace pcp {
// the unpruned kernels; dist_ws: b * n floats for the streaming fallback (nullptr: library-owned, grown with cudaMalloc)
static int fps_plain(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, float *dist_ws, cudaStream_t st)
{
    // smallest cluster whose register file holds the cloud: cs * 512 threads * PPT points, PPT <= 25.
    // Prefer the widest PPT that fits (fewest CTAs per cloud => most clouds in flight), but widen the
    // cluster while the grid would leave SMs idle.
    constexpr int kMaxPPT = 25;
    int cs = 1;
    while (cs < kFpsMaxCluster && (n > (int64_t)cs * kFpsThreads * kMaxPPT)) cs *= 2;
    while (cs < 8 && (int64_t)b * cs * 2 <= sm_count() && n / cs > 2 * kFpsThreads) cs *= 2;
    if (n > (int64_t)cs * kFpsThreads * kMaxPPT) {
        float *dist = dist_ws;
        if (!dist) {      // no caller workspace (pcp_fps_f32): per-device scratch, grown on demand — calls that overlap on the GPU share it
            const int dev = fps_device();
            PCP_REQUIRE(dev >= 0, PCP_ERR_CUDA, "fps: no current device");
            const size_t need = (size_t)b * (size_t)n * sizeof(float);
            if (need > g_fps_scratch_bytes[dev]) {
                if (g_fps_scratch[dev]) PCP_CUDA_OK(cudaFree(g_fps_scratch[dev]));
                g_fps_scratch[dev] = nullptr; g_fps_scratch_bytes[dev] = 0;
                PCP_CUDA_OK(cudaMalloc(&g_fps_scratch[dev], need));
                g_fps_scratch_bytes[dev] = need;
            }
            dist = g_fps_scratch[dev];
        }
        fps_stream_kernel<<<b, 1024, 0, st>>>(d_x, n, s, start, d_idx, dist);
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
}  // namespace pcp

/* development aid: cycles block 0 / thread 0 of the last PCP_FPS_PROF=1 launch (PPT = 25) spent in: distance loop, candidate
 * search + record, __syncthreads, warp-0 stage + DSMEM pushes, cluster barrier, slot read-back */
extern "C" int pcp_fps_phase_cycles(unsigned long long out[8])
{
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpyFromSymbol(out, g_fps_phase, 8 * sizeof(unsigned long long)));
    return PCP_OK;
}
