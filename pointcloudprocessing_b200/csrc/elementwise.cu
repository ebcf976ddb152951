// elementwise.cu — per-cloud normalisation, jitter, rotation, resize-to-width, gather.
// HBM-bound streaming kernels: 24 B/point algorithmic traffic (12 read + 12 written).
//   pcp_normalize_f32      <- PointCloudNormalization.call   (pointnet/PointNet.py:691-706)
//   pcp_jitter_f32         <- PointCloudSet._parse_function  (pointcloud/PointCloudSet.py:335-337)
//   pcp_resize_to_width_f32<- _adjust_to_input_width         (pointcloud/PointCloudSet.py:443-470)
//   pcp_rotate_f32         <- new (DCM convention utils/mat_ops.py:6-39)
#include <cooperative_groups.h>
#include <algorithm>
#include <cstdlib>

#include "common.cuh"
#include "umma.cuh"

namespace cg = cooperative_groups;

namespace pcp {
using namespace umma;

constexpr int kNormThreads = 512;

// One thread-block CLUSTER per cloud.  Each CTA stages its slice of the cloud in shared memory
// with 128-bit loads (the only HBM read), the cluster reduces centroid (double) and max distance
// through distributed shared memory, and each CTA writes its slice back normalised.
__global__ void __launch_bounds__(kNormThreads)
normalize_cluster_kernel(const float *__restrict__ x, int64_t n, float *__restrict__ y,
                         float *__restrict__ centroid_out, float *__restrict__ scale_out, int cs, int64_t ppc)
{
    extern __shared__ float4 smem4[];
    float *sm = reinterpret_cast<float *>(smem4);
    __shared__ double red_d[kNormThreads / 32][3];
    __shared__ double part_sum[3];
    __shared__ float red_f[kNormThreads / 32];
    __shared__ float part_max;

    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int64_t cloud = blockIdx.x / cs;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const int64_t p0 = min(n, (int64_t)rank * ppc), p1 = min(n, p0 + ppc);
    const int64_t np = p1 - p0, nf = np * 3;
    const float *src = x + (cloud * n + p0) * 3;
    float *dst = y + (cloud * n + p0) * 3;
    // element i of the slice lives at sm[off + i] so that 16-byte global and shared addresses line up
    const int off = (int)((reinterpret_cast<uintptr_t>(src) & 15u) >> 2);
    const bool dst_same = ((reinterpret_cast<uintptr_t>(dst) & 15u) >> 2) == (unsigned)off;

    const int64_t nvec = (off + nf + 3) >> 2;
    __shared__ uint64_t load_bar;
    const bool bulk = off == 0 && ((nf * 4) & 15) == 0 && nf > 0;      // CTA-uniform
    if (bulk) {
        // the slice is 16-byte aligned and sized: ONE thread streams it into shared memory with bulk async copies (every
        // 32 KB chunk in flight at once) — a per-thread float4 loop keeps only one 16-byte load per thread outstanding
        // (8 KB per SM), which capped this kernel at ~18 % of the HBM roofline
        if (tid == 0) { mbar_init(&load_bar, 1); fence_barrier_init(); }
        __syncthreads();
        if (tid == 0) {
            const uint32_t bytes = (uint32_t)(nf * 4);
            mbar_arrive_expect_tx(&load_bar, bytes);
            for (uint32_t o = 0; o < bytes; o += 32768u)
                bulk_g2s(reinterpret_cast<char *>(sm) + o, reinterpret_cast<const char *>(src) + o, min(32768u, bytes - o), &load_bar);
        }
        mbar_wait(&load_bar, 0, 70);
    } else {
        for (int64_t v0 = tid; v0 < nvec; v0 += 4 * kNormThreads) {
            float4 q[4];
            bool full[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t v = v0 + (int64_t)u * kNormThreads, base = 4 * v - off;
                full[u] = v < nvec && base >= 0 && base + 3 < nf;
                if (full[u]) q[u] = __ldg(reinterpret_cast<const float4 *>(src + base));
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int64_t v = v0 + (int64_t)u * kNormThreads, base = 4 * v - off;
                if (full[u]) {
                    smem4[v] = q[u];
                } else if (v < nvec) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (base + j >= 0 && base + j < nf) sm[4 * v + j] = __ldg(src + base + j);
                }
            }
        }
    }
    __syncthreads();

    // ---- centroid: double accumulation => correctly rounded mean (oracle/pointnet_numpy.py:normalize) ----
    double sx = 0, sy = 0, sz = 0;
    for (int64_t i = tid; i < np; i += kNormThreads) {
        sx += (double)sm[off + 3 * i];
        sy += (double)sm[off + 3 * i + 1];
        sz += (double)sm[off + 3 * i + 2];
    }
    sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
    if (lane == 0) { red_d[warp][0] = sx; red_d[warp][1] = sy; red_d[warp][2] = sz; }
    __syncthreads();
    if (tid < 3) {
        double s = 0;
        for (int w = 0; w < kNormThreads / 32; ++w) s += red_d[w][tid];
        part_sum[tid] = s;
    }
    cluster.sync();
    double tot[3] = {0, 0, 0};
    for (int r = 0; r < cs; ++r) {
        const double *rp = cluster.map_shared_rank(part_sum, r);
        tot[0] += rp[0]; tot[1] += rp[1]; tot[2] += rp[2];
    }
    const float cx = (float)(tot[0] / (double)n), cy = (float)(tot[1] / (double)n), cz = (float)(tot[2] / (double)n);

    // ---- max squared distance: (dx*dx + dy*dy) + dz*dz, separate roundings ----
    float md = 0.f;
    for (int64_t i = tid; i < np; i += kNormThreads) {
        const float dx = __fsub_rn(sm[off + 3 * i], cx);
        const float dy = __fsub_rn(sm[off + 3 * i + 1], cy);
        const float dz = __fsub_rn(sm[off + 3 * i + 2], cz);
        const float dd = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
        md = fmaxf(md, dd);
    }
    md = warp_max(md);
    if (lane == 0) red_f[warp] = md;
    __syncthreads();
    if (tid == 0) {
        float m = 0.f;
        for (int w = 0; w < kNormThreads / 32; ++w) m = fmaxf(m, red_f[w]);
        part_max = m;
    }
    cluster.sync();
    float gmax = 0.f;
    for (int r = 0; r < cs; ++r) gmax = fmaxf(gmax, *cluster.map_shared_rank(&part_max, r));
    const float scale = fmaxf(__fsqrt_rn(gmax), 1e-7f);   // sqrt is monotone: sqrt(max d^2) == max sqrt(d^2)
    if (rank == 0 && tid == 0) {
        if (centroid_out) { centroid_out[cloud * 3] = cx; centroid_out[cloud * 3 + 1] = cy; centroid_out[cloud * 3 + 2] = cz; }
        if (scale_out) scale_out[cloud] = scale;
    }

    // ---- write back: (x - c) / scale, true divide ----
    const float c3[3] = {cx, cy, cz};
    if (y == nullptr) {
        // statistics only (the fused model path normalises inside its first chain kernel)
    } else if (dst_same) {
        for (int64_t v = tid; v < nvec; v += kNormThreads) {
            const int64_t base = 4 * v - off;
            // (p0*3 + base) % 3 == base % 3 (mod with negatives handled below)
            int comp = (int)(((base % 3) + 3) % 3);
            if (base >= 0 && base + 3 < nf) {
                float4 q = smem4[v];
                float r[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    r[j] = __fdiv_rn(__fsub_rn(r[j], c3[comp]), scale);
                    comp = comp == 2 ? 0 : comp + 1;
                }
                *reinterpret_cast<float4 *>(dst + base) = make_float4(r[0], r[1], r[2], r[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (base + j >= 0 && base + j < nf) dst[base + j] = __fdiv_rn(__fsub_rn(sm[4 * v + j], c3[comp]), scale);
                    comp = comp == 2 ? 0 : comp + 1;
                }
            }
        }
    } else {
        for (int64_t i = tid; i < nf; i += kNormThreads) dst[i] = __fdiv_rn(__fsub_rn(sm[off + i], c3[i % 3]), scale);
    }
    cluster.sync();   // keep part_sum / part_max alive until every peer has read them
}

// Statistics only (centroid, scale) for batches of SMALL clouds — what the fused model path needs before its first chain
// kernel.  A cluster launch per 12 KB cloud is all launch / barrier overhead (74 us for 4096 x 1024 points); here a warp
// owns a cloud: fp64 sums, then the max squared distance on the second (L1-resident) pass.  Same arithmetic as
// normalize_cluster_kernel.
__global__ void __launch_bounds__(256)
normalize_stats_warp_kernel(const float *__restrict__ x, int b, int64_t n, float *__restrict__ centroid_out, float *__restrict__ scale_out)
{
    const int lane = threadIdx.x & 31;
    const int64_t cloud = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (cloud >= b) return;
    const float *xc = x + cloud * n * 3;
    double sx = 0, sy = 0, sz = 0;
    for (int64_t i = lane; i < n; i += 32) { sx += (double)xc[3 * i]; sy += (double)xc[3 * i + 1]; sz += (double)xc[3 * i + 2]; }
    sx = warp_sum(sx); sy = warp_sum(sy); sz = warp_sum(sz);
    sx = __shfl_sync(0xffffffffu, sx, 0); sy = __shfl_sync(0xffffffffu, sy, 0); sz = __shfl_sync(0xffffffffu, sz, 0);
    const float cx = (float)(sx / (double)n), cy = (float)(sy / (double)n), cz = (float)(sz / (double)n);
    float md = 0.f;
    for (int64_t i = lane; i < n; i += 32) {
        const float dx = __fsub_rn(xc[3 * i], cx), dy = __fsub_rn(xc[3 * i + 1], cy), dz = __fsub_rn(xc[3 * i + 2], cz);
        md = fmaxf(md, __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    }
    md = warp_max(md);
    if (lane == 0) {
        centroid_out[cloud * 3] = cx; centroid_out[cloud * 3 + 1] = cy; centroid_out[cloud * 3 + 2] = cz;
        scale_out[cloud] = fmaxf(__fsqrt_rn(md), 1e-7f);
    }
}

// ---- streaming kernels: 4 points (3 x float4) per thread-iteration ----
__device__ __forceinline__ void box_muller3(const uint32_t r[4], float out[3]) {
    // u in (0,1]: (r + 1) * 2^-32 ; normals: sqrt(-2 ln u0) * {cos,sin}(2 pi u1), sqrt(-2 ln u2) * cos(2 pi u3)
    const float u0 = ((float)(r[0] >> 8) + 1.0f) * (1.0f / 16777216.0f);
    const float u1 = (float)(r[1] >> 8) * (1.0f / 16777216.0f);
    const float u2 = ((float)(r[2] >> 8) + 1.0f) * (1.0f / 16777216.0f);
    const float u3 = (float)(r[3] >> 8) * (1.0f / 16777216.0f);
    const float ra = sqrtf(-2.0f * logf(u0)), rb = sqrtf(-2.0f * logf(u2));
    float s, c;
    sincospif(2.0f * u1, &s, &c);
    out[0] = ra * c; out[1] = ra * s;
    out[2] = rb * cospif(2.0f * u3);
}

__device__ __forceinline__ void point_noise(int64_t point, uint64_t seed, uint64_t offset, float out[3]) {
    const uint64_t ctr = (uint64_t)point + offset;
    uint32_t c[4] = {(uint32_t)ctr, (uint32_t)(ctr >> 32), 0u, 0u};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    box_muller3(c, out);
}

// Warp-cooperative access to runs of xyz points with fully coalesced 16-byte transactions.  A lane that owns 4 points owns
// 48 contiguous bytes, so per-lane float4 accesses touch every 32-byte sector from two different instructions (half-used
// sectors, twice the L1 wavefronts and L2 store transactions).  Instead lane l moves float4 l, l+32, l+64 of the warp's
// 1536-byte run and the lanes pick up / drop off their own 48 bytes in a warp-private shared tile (conflict-free: the
// quarter-warps of a 16-byte shared access at a 48-byte stride cover all 32 banks).
struct WarpRun {
    float4 *tile;                       // 96 float4, private to the warp
    int lane;
    __device__ void load(const float4 *src, float4 (&q)[3]) const { q[0] = __ldg(src + lane); q[1] = __ldg(src + 32 + lane); q[2] = __ldg(src + 64 + lane); }
    __device__ void scatter(const float4 (&q)[3], float (&v)[12]) const
    {
        tile[lane] = q[0]; tile[32 + lane] = q[1]; tile[64 + lane] = q[2];
        __syncwarp();
        const float4 a = tile[3 * lane], b = tile[3 * lane + 1], c = tile[3 * lane + 2];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w; v[8] = c.x; v[9] = c.y; v[10] = c.z; v[11] = c.w;
        __syncwarp();
    }
    __device__ void store(const float (&v)[12], float4 *dst) const
    {
        tile[3 * lane] = make_float4(v[0], v[1], v[2], v[3]);
        tile[3 * lane + 1] = make_float4(v[4], v[5], v[6], v[7]);
        tile[3 * lane + 2] = make_float4(v[8], v[9], v[10], v[11]);
        __syncwarp();
        dst[lane] = tile[lane]; dst[32 + lane] = tile[32 + lane]; dst[64 + lane] = tile[64 + lane];
        __syncwarp();
    }
};

// ---- normalisation of LiDAR-sized clouds (n >= kNormPartsMin): persistent CTAs, slices resident across the phases ----
// normalize_cluster_kernel keeps a cloud on chip too, but a cluster needs eight free SMs of one GPC at once and holds one
// CTA per SM that cannot do anything before its 150-200 KB load has landed: 27 % of the HBM roofline at 64 x 100 000.
// Here a cloud is cut into k slices of <= 48 KB; four CTAs per SM take (cloud, slice) tickets in order, keep their slice
// in shared memory (one bulk async copy = the only HBM read) and meet the cloud's other slices through two counters in
// global memory: fp64 component sums -> arrive -> centroid; max squared distance -> arrive -> scale; write.  Tickets are
// issued in increasing order, so every slice a CTA can wait for is already held by a running CTA or is the next ticket
// (needs k resident CTAs, nothing about dispatch order).  Same arithmetic as normalize_cluster_kernel, so the same bits.
// What bounds it (scripts/norm_tune.py, PCP_NORM_DBG): a slice lives ~15 us in shared memory (load ~5, two arrive / poll /
// read-back chains of ~2.5 us of L2 round trips each, compute ~2, write ~2), so throughput = shared memory per SM / slice
// lifetime ~ 2.2 TB/s of input whatever the slice size; a second buffer per CTA or a static slice order did not help.
struct NormScratch { double sum[3]; unsigned maxd2, arrive1, arrive2, pad[3]; };   // 48 B per cloud, zeroed by the launcher
constexpr int64_t kNormPartsMin = 16384;
constexpr int kNormSliceBytes = 48 * 1024;       // four 256-thread CTAs per SM
constexpr int kNormPartThreads = 256;

__device__ __forceinline__ void norm_spin(const unsigned *flag, unsigned target)
{
    const long long t0 = clock64();
    unsigned v;
    do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v >= target) break;
        __nanosleep(64);
        if (clock64() - t0 > 4000000000ll) __trap();        // a lost arrival must not hang the device
    } while (true);
}

// (x - c) / s correctly rounded without the MUFU.RCP + Newton sequence per element: with r = RN(1 / s), q0 = RN(a r),
// e = a - s q0 (exact, one FMA), q = RN(q0 + e r) is the correctly rounded quotient (Markstein) unless s has an all-ones
// significand or an intermediate leaves the normal range — those take the IEEE divide.
struct NormDiv {
    float s, r;
    bool fast;
    __device__ __forceinline__ explicit NormDiv(float scale)
    {
        s = scale;
        r = __frcp_rn(scale);
        const unsigned u = __float_as_uint(scale);
        fast = (u & 0x007fffffu) != 0x007fffffu && scale > 1e-30f && scale < 1e30f;
    }
    __device__ __forceinline__ float operator()(float a) const
    {
        const float q0 = __fmul_rn(a, r);
        const float q = __fmaf_rn(__fmaf_rn(-q0, s, a), r, q0);
        return (fast && fabsf(q0) > 1e-30f && fabsf(q0) < 1e30f) ? q : __fdiv_rn(a, s);
    }
};

__global__ void __launch_bounds__(kNormPartThreads)
normalize_parts_kernel(const float *__restrict__ x, int b, int64_t n, int k, int64_t ppc, float *__restrict__ y,
                       float *__restrict__ centroid_out, float *__restrict__ scale_out, NormScratch *__restrict__ scr,
                       unsigned *__restrict__ ticket, unsigned dbg)
{   // (dbg: timing-experiment bits, see scripts/norm_tune.py; 0 in production)
    constexpr int T = kNormPartThreads;                    // T % 3 == 1: float4 tid + it T of the slice starts at axis (tid + it) % 3
    static_assert(T % 3 == 1, "axis rotation below assumes T = 1 (mod 3)");
    extern __shared__ float4 smem4[];
    __shared__ double red_d[T / 32][3];
    __shared__ float red_f[T / 32];
    __shared__ uint64_t load_bar;
    __shared__ unsigned item_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, r0 = tid % 3;
    const unsigned total = (unsigned)b * (unsigned)k;
    if (tid == 0) { mbar_init(&load_bar, 1); fence_barrier_init(); }
    uint32_t parity = 0;
    while (true) {
        __syncthreads();                                   // the previous slice is dead (and load_bar is initialised)
        if (tid == 0) item_s = atomicAdd(ticket, 1u);
        __syncthreads();
        const unsigned item = item_s;
        if (item >= total) break;
        const int64_t cloud = item / (unsigned)k;
        const int64_t p0 = min(n, (int64_t)(item % (unsigned)k) * ppc);
        const int np = (int)(min(n, p0 + ppc) - p0);       // np % 4 == 0
        const int nvec = (np * 3) >> 2;
        NormScratch *sc = scr + cloud;
        if (np > 0) {
            if (tid == 0) {
                const uint32_t bytes = (uint32_t)np * 12u;
                const char *src = reinterpret_cast<const char *>(x + (cloud * n + p0) * 3);
                mbar_arrive_expect_tx(&load_bar, bytes);
                for (uint32_t o = 0; o < bytes; o += 32768u)
                    bulk_g2s(reinterpret_cast<char *>(smem4) + o, src + o, min(32768u, bytes - o), &load_bar);
            }
            mbar_wait(&load_bar, parity, 71);
            parity ^= 1u;
        }
        // ---- fp64 component sums: 16-byte reads, accumulator [u][j] holds axis (r0 + u + j) % 3 ----
        double acc[3][4];
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[u][j] = 0.0;
        for (int v0 = tid; v0 < ((dbg & 2) ? 0 : nvec); v0 += 3 * T) {
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                if (v0 + u * T < nvec) {
                    const float4 q = smem4[v0 + u * T];
                    acc[u][0] += (double)q.x; acc[u][1] += (double)q.y; acc[u][2] += (double)q.z; acc[u][3] += (double)q.w;
                }
            }
        }
        double s3[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int axis = (r0 + u + j) % 3;
                s3[0] += axis == 0 ? acc[u][j] : 0.0; s3[1] += axis == 1 ? acc[u][j] : 0.0; s3[2] += axis == 2 ? acc[u][j] : 0.0;
            }
#pragma unroll
        for (int a = 0; a < 3; ++a) { s3[a] = warp_sum(s3[a]); if (lane == 0) red_d[warp][a] = s3[a]; }
        __syncthreads();
        if (tid < 3) {
            double t = 0;
            for (int w = 0; w < T / 32; ++w) t += red_d[w][tid];
            atomicAdd(&sc->sum[tid], t);
            __threadfence();
        }
        __syncthreads();
        if (tid == 0) { atomicAdd(&sc->arrive1, 1u); if (!(dbg & 1)) norm_spin(&sc->arrive1, (unsigned)k); }
        __syncthreads();
        const float cx = (float)(__ldcg(&sc->sum[0]) / (double)n), cy = (float)(__ldcg(&sc->sum[1]) / (double)n),
                    cz = (float)(__ldcg(&sc->sum[2]) / (double)n);
        // ---- max squared distance: (dx*dx + dy*dy) + dz*dz, separate roundings; four points = three 16-byte reads ----
        float md = 0.f;
        auto dist = [&](float px, float py, float pz) {
            const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy), dz = __fsub_rn(pz, cz);
            md = fmaxf(md, __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
        };
        for (int g = tid; g < ((dbg & 2) ? 0 : (np >> 2)); g += T) {
            const float4 q0 = smem4[3 * g], q1 = smem4[3 * g + 1], q2 = smem4[3 * g + 2];
            dist(q0.x, q0.y, q0.z); dist(q0.w, q1.x, q1.y); dist(q1.z, q1.w, q2.x); dist(q2.y, q2.z, q2.w);
        }
        md = warp_max(md);
        if (lane == 0) red_f[warp] = md;
        __syncthreads();
        if (tid == 0) {
            float m = 0.f;
            for (int w = 0; w < T / 32; ++w) m = fmaxf(m, red_f[w]);
            atomicMax(&sc->maxd2, __float_as_uint(m));      // non-negative floats order as integers
            __threadfence();
            atomicAdd(&sc->arrive2, 1u);
            if (!(dbg & 1)) norm_spin(&sc->arrive2, (unsigned)k);
        }
        __syncthreads();
        const float scale = fmaxf(__fsqrt_rn(__uint_as_float(__ldcg(&sc->maxd2))), 1e-7f);
        if (p0 == 0 && tid == 0) {
            if (centroid_out) { centroid_out[cloud * 3] = cx; centroid_out[cloud * 3 + 1] = cy; centroid_out[cloud * 3 + 2] = cz; }
            if (scale_out) scale_out[cloud] = scale;
        }
        // ---- write back: (x - c) / scale, true divide ----
        if (y) {
            float4 *dst = reinterpret_cast<float4 *>(y + (cloud * n + p0) * 3);
            const NormDiv div(scale);
            const float c3[3] = {cx, cy, cz};
            float cu[3][3];                                 // centroid components in the order float4 (tid + (3 m + u) T) sees them
#pragma unroll
            for (int u = 0; u < 3; ++u)
#pragma unroll
                for (int j = 0; j < 3; ++j) {
                    const int axis = (r0 + u + j) % 3;
                    cu[u][j] = axis == 0 ? c3[0] : axis == 1 ? c3[1] : c3[2];
                }
            for (int v0 = tid; v0 < ((dbg & 4) ? 0 : nvec); v0 += 3 * T) {
#pragma unroll
                for (int u = 0; u < 3; ++u) {
                    if (v0 + u * T < nvec) {
                        const float4 q = smem4[v0 + u * T];
                        dst[v0 + u * T] = make_float4(div(__fsub_rn(q.x, cu[u][0])), div(__fsub_rn(q.y, cu[u][1])), div(__fsub_rn(q.z, cu[u][2])),
                                                      div(__fsub_rn(q.w, cu[u][0])));
                    }
                }
            }
        }
    }
}

__global__ void __launch_bounds__(256)
philox_normal_kernel(int64_t npoints, uint64_t seed, uint64_t offset, float *__restrict__ out)
{
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < npoints; p += (int64_t)gridDim.x * blockDim.x) {
        float z[3];
        point_noise(p, seed, offset, z);
        out[3 * p] = z[0]; out[3 * p + 1] = z[1]; out[3 * p + 2] = z[2];
    }
}

// y = x + noise * stdev (fp32 multiply, then add — PointCloudSet.py:337 is not fused)
template <bool kExplicitNoise>
__global__ void __launch_bounds__(256)
jitter_kernel(const float *__restrict__ x, const float *__restrict__ noise, int64_t npoints, float sx, float sy, float sz,
              uint64_t seed, uint64_t offset, float *__restrict__ y, int vec_ok)
{
    const float sd[3] = {sx, sy, sz};
    const int64_t ngroups = npoints >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    __shared__ float4 tiles[8][96];
    if (vec_ok) {
        // full warp runs (128 points): coalesced through the warp tile, two runs in flight per warp
        // (the Philox variant is instruction bound — 10 Philox rounds + Box-Muller per point — and keeps the direct path)
        const WarpRun wr{tiles[threadIdx.x >> 5], (int)(threadIdx.x & 31)};
        const int64_t nruns = kExplicitNoise ? (ngroups >> 5) : 0, nwarps = stride >> 5, w0 = t0 >> 5;
        const float4 *x4 = reinterpret_cast<const float4 *>(x), *n4 = reinterpret_cast<const float4 *>(noise);
        float4 *y4 = reinterpret_cast<float4 *>(y);
        auto body = [&](int64_t run, const float4 (&q)[3], const float4 (&qn)[3]) {
            float v[12], z[12];
            wr.scatter(q, v);
            if (kExplicitNoise) {
                wr.scatter(qn, z);
            } else {
                const int64_t g = run * 32 + wr.lane;
#pragma unroll
                for (int p = 0; p < 4; ++p) point_noise(4 * g + p, seed, offset, z + 3 * p);
            }
#pragma unroll
            for (int j = 0; j < 12; ++j) v[j] = __fadd_rn(v[j], __fmul_rn(z[j], sd[j % 3]));
            wr.store(v, y4 + run * 96);
        };
        for (int64_t run = w0; run < nruns; run += 2 * nwarps) {
            float4 q0[3], q1[3], m0[3], m1[3];
            const bool two = run + nwarps < nruns;
            wr.load(x4 + run * 96, q0);
            if (kExplicitNoise) wr.load(n4 + run * 96, m0);
            if (two) {
                wr.load(x4 + (run + nwarps) * 96, q1);
                if (kExplicitNoise) wr.load(n4 + (run + nwarps) * 96, m1);
            }
            body(run, q0, m0);
            if (two) body(run + nwarps, q1, m1);
        }
        for (int64_t g = (nruns << 5) + t0; g < ngroups; g += stride) {       // the last partial run
            const float4 *xs = reinterpret_cast<const float4 *>(x) + 3 * g;
            float4 a = __ldg(xs), b = __ldg(xs + 1), c = __ldg(xs + 2);
            float v[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
            float z[12];
            if (kExplicitNoise) {
                const float4 *ns = reinterpret_cast<const float4 *>(noise) + 3 * g;
                float4 na = __ldg(ns), nb = __ldg(ns + 1), nc = __ldg(ns + 2);
                z[0] = na.x; z[1] = na.y; z[2] = na.z; z[3] = na.w; z[4] = nb.x; z[5] = nb.y; z[6] = nb.z; z[7] = nb.w;
                z[8] = nc.x; z[9] = nc.y; z[10] = nc.z; z[11] = nc.w;
            } else {
#pragma unroll
                for (int p = 0; p < 4; ++p) point_noise(4 * g + p, seed, offset, z + 3 * p);
            }
#pragma unroll
            for (int j = 0; j < 12; ++j) v[j] = __fadd_rn(v[j], __fmul_rn(z[j], sd[j % 3]));
            float4 *ys = reinterpret_cast<float4 *>(y) + 3 * g;
            ys[0] = make_float4(v[0], v[1], v[2], v[3]);
            ys[1] = make_float4(v[4], v[5], v[6], v[7]);
            ys[2] = make_float4(v[8], v[9], v[10], v[11]);
        }
    }
    const int64_t tail0 = vec_ok ? (ngroups << 2) : 0;
    for (int64_t p = tail0 + t0; p < npoints; p += stride) {
        float z[3];
        if (kExplicitNoise) { z[0] = noise[3 * p]; z[1] = noise[3 * p + 1]; z[2] = noise[3 * p + 2]; }
        else point_noise(p, seed, offset, z);
#pragma unroll
        for (int j = 0; j < 3; ++j) y[3 * p + j] = __fadd_rn(x[3 * p + j], __fmul_rn(z[j], sd[j]));
    }
}

// y_j = (x0*R[j][0] + x1*R[j][1]) + x2*R[j][2]  (no FMA), R per cloud.  grid.y = cloud.
__global__ void __launch_bounds__(256)
rotate_kernel(const float *__restrict__ x, int64_t n, const float *__restrict__ R, float *__restrict__ y, int vec_ok)
{
    const int64_t cloud = blockIdx.y;
    __shared__ float r[9];
    if (threadIdx.x < 9) r[threadIdx.x] = R[cloud * 9 + threadIdx.x];
    __syncthreads();
    const float *xc = x + cloud * n * 3;
    float *yc = y + cloud * n * 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t ngroups = n >> 2;
    auto rot = [&](const float *p, float *q) {
#pragma unroll
        for (int j = 0; j < 3; ++j)
            q[j] = __fadd_rn(__fadd_rn(__fmul_rn(p[0], r[3 * j]), __fmul_rn(p[1], r[3 * j + 1])), __fmul_rn(p[2], r[3 * j + 2]));
    };
    __shared__ float4 tiles[8][96];
    if (vec_ok) {
        const WarpRun wr{tiles[threadIdx.x >> 5], (int)(threadIdx.x & 31)};
        const int64_t nruns = ngroups >> 5, nwarps = stride >> 5, w0 = t0 >> 5;
        const float4 *x4 = reinterpret_cast<const float4 *>(xc);
        float4 *y4 = reinterpret_cast<float4 *>(yc);
        auto body = [&](int64_t run, const float4 (&q)[3]) {
            float v[12], o[12];
            wr.scatter(q, v);
#pragma unroll
            for (int p = 0; p < 4; ++p) rot(v + 3 * p, o + 3 * p);
            wr.store(o, y4 + run * 96);
        };
        for (int64_t run = w0; run < nruns; run += 2 * nwarps) {
            float4 q0[3], q1[3];
            const bool two = run + nwarps < nruns;
            wr.load(x4 + run * 96, q0);
            if (two) wr.load(x4 + (run + nwarps) * 96, q1);
            body(run, q0);
            if (two) body(run + nwarps, q1);
        }
        for (int64_t g = (nruns << 5) + t0; g < ngroups; g += stride) {       // the last partial run
            const float4 *xs = reinterpret_cast<const float4 *>(xc) + 3 * g;
            float4 a = __ldg(xs), b = __ldg(xs + 1), c = __ldg(xs + 2);
            float v[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
            float o[12];
#pragma unroll
            for (int p = 0; p < 4; ++p) rot(v + 3 * p, o + 3 * p);
            float4 *ys = reinterpret_cast<float4 *>(yc) + 3 * g;
            ys[0] = make_float4(o[0], o[1], o[2], o[3]);
            ys[1] = make_float4(o[4], o[5], o[6], o[7]);
            ys[2] = make_float4(o[8], o[9], o[10], o[11]);
        }
    }
    const int64_t tail0 = vec_ok ? (ngroups << 2) : 0;
    for (int64_t p = tail0 + t0; p < n; p += stride) {
        float v[3] = {xc[3 * p], xc[3 * p + 1], xc[3 * p + 2]}, o[3];
        rot(v, o);
        yc[3 * p] = o[0]; yc[3 * p + 1] = o[1]; yc[3 * p + 2] = o[2];
    }
}

// _adjust_to_input_width: slot j < N -> point j; slot j >= N -> point floor(u * N), u from Philox(seed; cloud, j)
__global__ void __launch_bounds__(256)
resize_kernel(const float *__restrict__ x, const int32_t *__restrict__ labels, int64_t n_in_stride,
              const int32_t *__restrict__ n_in, int64_t n_out, uint64_t seed, float *__restrict__ y,
              int32_t *__restrict__ labels_out)
{
    const int64_t cloud = blockIdx.y;
    const int64_t N = n_in[cloud];
    const float *xc = x + cloud * n_in_stride * 3;
    float *yc = y + cloud * n_out * 3;
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < n_out; j += (int64_t)gridDim.x * blockDim.x) {
        int64_t srcp = j;
        if (j >= N) {
            uint32_t c[4] = {(uint32_t)j, (uint32_t)(j >> 32), (uint32_t)cloud, 0x5e512eu};
            philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
            // floor(U(0,N)) with U = r * 2^-32 evaluated exactly in integers
            srcp = N > 0 ? (int64_t)(((uint64_t)c[0] * (uint64_t)N) >> 32) : 0;
        }
        float vx = 0.f, vy = 0.f, vz = 0.f;
        int32_t lb = 0;
        if (N > 0) {
            vx = xc[3 * srcp]; vy = xc[3 * srcp + 1]; vz = xc[3 * srcp + 2];
            if (labels) lb = labels[cloud * n_in_stride + srcp];
        }
        yc[3 * j] = vx; yc[3 * j + 1] = vy; yc[3 * j + 2] = vz;
        if (labels_out) labels_out[cloud * n_out + j] = lb;
    }
}

__global__ void __launch_bounds__(256)
gather_kernel(const float *__restrict__ x, const int32_t *__restrict__ labels, int64_t n, const int32_t *__restrict__ idx,
              int s, float *__restrict__ out, int32_t *__restrict__ labels_out)
{
    const int64_t cloud = blockIdx.y;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < s; j += gridDim.x * blockDim.x) {
        const int64_t i = idx[cloud * s + j];
        const float *p = x + (cloud * n + i) * 3;
        float *q = out + (cloud * s + j) * 3;
        q[0] = p[0]; q[1] = p[1]; q[2] = p[2];
        if (labels_out) labels_out[cloud * s + j] = labels[cloud * n + i];
    }
}

static int stream_grid(int64_t work_items, int threads) {
    int64_t blocks = (work_items + threads - 1) / threads;
    const int64_t cap = (int64_t)sm_count() * 8;          // multiple of the SM count, grid-stride beyond
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

// ---- the sliced kernel's launch, shared by the three entry points that differ only in where the scratch lives ----
struct NormPartsCfg { bool ready = false; int slice_bytes = 0, ctas_per_sm = 0, dbg = 0; };
static NormPartsCfg g_norm_parts_cfg[64];            // per DEVICE: the shared-memory opt-in and the occupancy belong to a context

static int norm_parts_cfg(const NormPartsCfg **out)
{
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    PCP_REQUIRE(dev >= 0 && dev < 64, PCP_ERR_UNSUPPORTED, "normalize: device ordinal %d", dev);
    NormPartsCfg &c = g_norm_parts_cfg[dev];
    if (!c.ready) {
        c.slice_bytes = kNormSliceBytes;
        if (const char *e = getenv("PCP_NORM_SLICE_KB")) c.slice_bytes = std::max(4, std::min(200, atoi(e))) * 1024;   // tuning knob
        PCP_CUDA_OK(cudaFuncSetAttribute(normalize_parts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, c.slice_bytes + 384));   // (+ rounding of ppc)
        PCP_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c.ctas_per_sm, normalize_parts_kernel, kNormPartThreads, c.slice_bytes + 384));
        PCP_REQUIRE(c.ctas_per_sm >= 1, PCP_ERR_CUDA, "normalize: slice kernel does not fit an SM");
        if (const char *e = getenv("PCP_NORM_DBG")) c.dbg = atoi(e) & 7;
        c.ready = true;
    }
    *out = &c;
    return PCP_OK;
}

static inline bool norm_parts_applies(const float *d_x, const float *d_y, int b, int64_t n)
{
    const bool aligned = (((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y)) & 15u) == 0) && (n % 4 == 0);
    return n >= kNormPartsMin && aligned && b <= 65535;
}
static inline size_t norm_parts_scratch_bytes(int b) { return 64 + (size_t)b * sizeof(NormScratch); }

// slot: 64 bytes of ticket + b NormScratch records, cleared here on the stream
static int norm_parts_launch(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale, void *slot_v, cudaStream_t st)
{
    const NormPartsCfg *c = nullptr;
    PCP_TRY(norm_parts_cfg(&c));
    char *slot = static_cast<char *>(slot_v);
    PCP_CUDA_OK(cudaMemsetAsync(slot, 0, norm_parts_scratch_bytes(b), st));
    const int k = (int)((n * 12 + c->slice_bytes - 1) / c->slice_bytes);
    // slices start on 128-byte lines when the cloud does (32 points = 384 B): bulk copies of misaligned slices split sectors
    const int64_t ppc = (((n + k - 1) / k) + 31) & ~(int64_t)31;
    const int grid = (int)std::min<int64_t>((int64_t)b * k, (int64_t)sm_count() * c->ctas_per_sm);
    PCP_REQUIRE(k <= grid, PCP_ERR_UNSUPPORTED, "normalize: n=%lld needs %d resident slices", (long long)n, k);
    normalize_parts_kernel<<<grid, kNormPartThreads, (size_t)ppc * 12, st>>>(d_x, b, n, k, ppc, d_y, d_centroid, d_scale,
                                                                           reinterpret_cast<NormScratch *>(slot + 64), reinterpret_cast<unsigned *>(slot),
                                                                           (unsigned)c->dbg);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

static int norm_small_launch(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale, cudaStream_t st);

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_normalize_f32(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale,
                                 pcp_stream stream)
{
    PCP_REQUIRE(d_x && (d_y || (d_centroid && d_scale)) && b >= 0 && n >= 1, PCP_ERR_INVALID, "pcp_normalize_f32: bad arguments (b=%d n=%lld)", b, (long long)n);
    if (b == 0) return PCP_OK;
    cudaStream_t st = as_stream(stream);
    if (!norm_parts_applies(d_x, d_y, b, n) || (!d_y && n <= 4096 && b >= 64)) return norm_small_launch(d_x, b, n, d_y, d_centroid, d_scale, st);
    // Library-owned scratch for callers that pass none (pcp_normalize_ws_f32 is the allocation-free form): a per-device ring of
    // eight slots; a slot is handed out again only after the stream that used it last has passed its kernel (event per slot), so
    // any number of calls may be in flight on any number of streams.  Growing the ring frees and allocates (device-synchronising).
    struct Ring { char *base = nullptr; int64_t cap = 0; unsigned turn = 0; cudaEvent_t done[8] = {}; bool used[8] = {}; };
    static Ring rings[64];
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    PCP_REQUIRE(dev >= 0 && dev < 64, PCP_ERR_UNSUPPORTED, "pcp_normalize_f32: device ordinal %d", dev);
    Ring &r = rings[dev];
    if (r.cap < b) {
        if (r.base) PCP_CUDA_OK(cudaFree(r.base));       // synchronises the device: no slot is in use afterwards
        r.base = nullptr;
        r.cap = std::max<int64_t>(b, 1024);
        PCP_CUDA_OK(cudaMalloc(&r.base, norm_parts_scratch_bytes((int)r.cap) * 8));
        for (bool &u : r.used) u = false;
    }
    const unsigned i = r.turn++ & 7u;
    if (!r.done[i]) PCP_CUDA_OK(cudaEventCreateWithFlags(&r.done[i], cudaEventDisableTiming));
    if (r.used[i]) PCP_CUDA_OK(cudaStreamWaitEvent(st, r.done[i], 0));
    PCP_TRY(norm_parts_launch(d_x, b, n, d_y, d_centroid, d_scale, r.base + (size_t)i * norm_parts_scratch_bytes((int)r.cap), st));
    PCP_CUDA_OK(cudaEventRecord(r.done[i], st));
    r.used[i] = true;
    return PCP_OK;
}

extern "C" int pcp_normalize_workspace_bytes(int b, int64_t n, size_t *bytes)
{
    PCP_REQUIRE(bytes && b >= 0 && n >= 1, PCP_ERR_INVALID, "pcp_normalize_workspace_bytes: bad arguments");
    // only the sliced kernel (LiDAR-sized clouds) uses scratch; alignment of x / y is the caller's to keep (16 bytes)
    *bytes = (b > 0 && n >= kNormPartsMin && n % 4 == 0 && b <= 65535) ? ((norm_parts_scratch_bytes(b) + 255) & ~(size_t)255) : 0;
    return PCP_OK;
}

extern "C" int pcp_normalize_ws_f32(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale, void *d_workspace,
                                    size_t workspace_bytes, pcp_stream stream)
{
    PCP_REQUIRE(d_x && (d_y || (d_centroid && d_scale)) && b >= 0 && n >= 1, PCP_ERR_INVALID, "pcp_normalize_ws_f32: bad arguments (b=%d n=%lld)", b,
                (long long)n);
    if (b == 0) return PCP_OK;
    cudaStream_t st = as_stream(stream);
    if (!norm_parts_applies(d_x, d_y, b, n) || (!d_y && n <= 4096 && b >= 64)) return norm_small_launch(d_x, b, n, d_y, d_centroid, d_scale, st);
    PCP_REQUIRE(d_workspace && workspace_bytes >= norm_parts_scratch_bytes(b) && (reinterpret_cast<uintptr_t>(d_workspace) & 15u) == 0, PCP_ERR_INVALID,
                "pcp_normalize_ws_f32: workspace of %zu bytes (16-byte aligned) needed, %zu given — size it with pcp_normalize_workspace_bytes",
                norm_parts_scratch_bytes(b), workspace_bytes);
    return norm_parts_launch(d_x, b, n, d_y, d_centroid, d_scale, d_workspace, st);
}

namespace pcp {
// clouds that need no scratch: a warp per cloud for the statistics-only pass of the model, else one thread-block cluster per cloud
static int norm_small_launch(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale, cudaStream_t st)
{
    if (!d_y && n <= 4096 && b >= 64) {
        normalize_stats_warp_kernel<<<(b + 7) / 8, 256, 0, st>>>(d_x, b, n, d_centroid, d_scale);
        PCP_CHECK_LAUNCH();
        return PCP_OK;
    }
    // cluster size: slice must fit in shared memory; widen further while the grid under-fills the GPU
    const int64_t max_slice_bytes = 200 * 1024;
    int cs = 1;
    while (cs < 16 && (n * 12 / cs > max_slice_bytes || ((int64_t)b * cs < 2 * sm_count() && n / cs > 4096))) cs *= 2;
    PCP_REQUIRE(n * 12 / cs <= max_slice_bytes, PCP_ERR_UNSUPPORTED,
                "pcp_normalize_f32: n=%lld exceeds the cluster-resident limit (%lld points)", (long long)n,
                (long long)(16 * max_slice_bytes / 12));
    const int64_t ppc = (((n + cs - 1) / cs) + 3) & ~(int64_t)3;
    const size_t smem = (size_t)(ppc * 3 + 8) * sizeof(float);
    static bool attr_set[64] = {};                     // per device
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    PCP_REQUIRE(dev >= 0 && dev < 64, PCP_ERR_UNSUPPORTED, "normalize: device ordinal %d", dev);
    if (!attr_set[dev]) {
        PCP_CUDA_OK(cudaFuncSetAttribute(normalize_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        PCP_CUDA_OK(cudaFuncSetAttribute(normalize_cluster_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr_set[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((int64_t)b * cs));
    cfg.blockDim = dim3(kNormThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, normalize_cluster_kernel, d_x, n, d_y, d_centroid, d_scale, cs, ppc));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}
}  // namespace pcp

extern "C" int pcp_philox_normal_f32(int64_t npoints, uint64_t seed, uint64_t offset, float *d_noise, pcp_stream stream)
{
    PCP_REQUIRE(d_noise && npoints >= 0, PCP_ERR_INVALID, "pcp_philox_normal_f32: bad arguments");
    if (npoints == 0) return PCP_OK;
    philox_normal_kernel<<<stream_grid(npoints, 256), 256, 0, as_stream(stream)>>>(npoints, seed, offset, d_noise);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_jitter_f32(const float *d_x, int64_t npoints, const float h_stdev[3], uint64_t seed, uint64_t offset,
                              const float *d_noise, float *d_y, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_y && h_stdev && npoints >= 0, PCP_ERR_INVALID, "pcp_jitter_f32: bad arguments");
    if (npoints == 0) return PCP_OK;
    const int vec_ok = ((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y) |
                         reinterpret_cast<uintptr_t>(d_noise)) & 15u) == 0;
    // one resident wave (grid-stride beyond): sm_count * 8 CTAs on a kernel that fits three or four per SM ends in a partial wave
    static int occ_explicit = 0, occ_philox = 0;
    if (!occ_explicit) {
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_explicit, jitter_kernel<true>, 256, 0) != cudaSuccess || occ_explicit < 1) occ_explicit = 4;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_philox, jitter_kernel<false>, 256, 0) != cudaSuccess || occ_philox < 1) occ_philox = 4;
    }
    const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(((npoints + 3) / 4 + 255) / 256, (int64_t)sm_count() * (d_noise ? occ_explicit : occ_philox)));
    if (d_noise)
        jitter_kernel<true><<<grid, 256, 0, as_stream(stream)>>>(d_x, d_noise, npoints, h_stdev[0], h_stdev[1], h_stdev[2], seed, offset, d_y, vec_ok);
    else
        jitter_kernel<false><<<grid, 256, 0, as_stream(stream)>>>(d_x, nullptr, npoints, h_stdev[0], h_stdev[1], h_stdev[2], seed, offset, d_y, vec_ok);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_rotate_f32(const float *d_x, int b, int64_t n, const float *d_R, float *d_y, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_y && d_R && b >= 0 && n >= 0, PCP_ERR_INVALID, "pcp_rotate_f32: bad arguments");
    if (b == 0 || n == 0) return PCP_OK;
    PCP_REQUIRE(b <= 65535, PCP_ERR_UNSUPPORTED, "pcp_rotate_f32: b > 65535");
    const int vec_ok = (((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y)) & 15u) == 0) && (n % 4 == 0);
    int gx = (int)((n / 4 + 255) / 256);
    const int cap = (sm_count() * 8 + b - 1) / b;
    if (gx > cap) gx = cap;
    if (gx < 1) gx = 1;
    rotate_kernel<<<dim3(gx, b), 256, 0, as_stream(stream)>>>(d_x, n, d_R, d_y, vec_ok);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_resize_to_width_f32(const float *d_x, const int32_t *d_labels, int b, int64_t n_in_stride,
                                       const int32_t *d_n_in, int64_t n_out, uint64_t seed, float *d_y,
                                       int32_t *d_labels_out, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_y && d_n_in && b >= 0 && n_out >= 0 && n_in_stride >= 0, PCP_ERR_INVALID, "pcp_resize_to_width_f32: bad arguments");
    PCP_REQUIRE(!d_labels_out || d_labels, PCP_ERR_INVALID, "pcp_resize_to_width_f32: labels_out without labels");
    if (b == 0 || n_out == 0) return PCP_OK;
    PCP_REQUIRE(b <= 65535, PCP_ERR_UNSUPPORTED, "pcp_resize_to_width_f32: b > 65535");
    int gx = (int)((n_out + 255) / 256);
    if (gx > 1024) gx = 1024;
    resize_kernel<<<dim3(gx, b), 256, 0, as_stream(stream)>>>(d_x, d_labels, n_in_stride, d_n_in, n_out, seed, d_y, d_labels_out);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_gather_points_f32(const float *d_x, const int32_t *d_labels, int b, int64_t n, const int32_t *d_idx, int s,
                                     float *d_out, int32_t *d_labels_out, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_idx && d_out && b >= 0 && s >= 0, PCP_ERR_INVALID, "pcp_gather_points_f32: bad arguments");
    PCP_REQUIRE(!d_labels_out || d_labels, PCP_ERR_INVALID, "pcp_gather_points_f32: labels_out without labels");
    if (b == 0 || s == 0) return PCP_OK;
    PCP_REQUIRE(b <= 65535, PCP_ERR_UNSUPPORTED, "pcp_gather_points_f32: b > 65535");
    gather_kernel<<<dim3((s + 255) / 256, b), 256, 0, as_stream(stream)>>>(d_x, d_labels, n, d_idx, s, d_out, d_labels_out);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// ---- allocation-free variant of the sliced normalisation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture, tight serving loops) ----
// Its scratch — the ticket word and one NormScratch per cloud — lives in a file-scope device array that every GPU which loads
// the library holds until exit.  The name `g_norm_noalloc_scratch` is only a host-side stand-in in host code: the device
// address is taken once per device with cudaGetSymbolAddress and that pointer is what the kernel receives.
#define PCP_NORMALIZE_NOALLOC_MAX_CLOUDS 4096
namespace pcp {
constexpr size_t kNormNoAllocBytes = 64 + (size_t)PCP_NORMALIZE_NOALLOC_MAX_CLOUDS * sizeof(NormScratch);      // 196 672 B
static __device__ __align__(256) unsigned char g_norm_noalloc_scratch[kNormNoAllocBytes];
}  // namespace pcp

extern "C" int pcp_normalize_noalloc_max_clouds(void) { return PCP_NORMALIZE_NOALLOC_MAX_CLOUDS; }

extern "C" int pcp_normalize_noalloc_f32(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale, pcp_stream stream)
{
    PCP_REQUIRE(d_x && (d_y || (d_centroid && d_scale)) && b >= 0 && n >= 1, PCP_ERR_INVALID, "pcp_normalize_noalloc_f32: bad arguments (b=%d n=%lld)", b,
                (long long)n);
    if (b == 0) return PCP_OK;
    PCP_REQUIRE(b <= PCP_NORMALIZE_NOALLOC_MAX_CLOUDS, PCP_ERR_UNSUPPORTED,
                "pcp_normalize_noalloc_f32: b=%d exceeds the static scratch (%d clouds); use pcp_normalize_f32", b, PCP_NORMALIZE_NOALLOC_MAX_CLOUDS);
    const bool aligned = (((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y)) & 15u) == 0) && (n % 4 == 0);
    PCP_REQUIRE(n >= kNormPartsMin && aligned, PCP_ERR_UNSUPPORTED,
                "pcp_normalize_noalloc_f32: only the sliced kernel uses scratch: n >= %lld, n %% 4 == 0 and 16-byte aligned x / y are required "
                "(smaller clouds need no scratch: pcp_normalize_f32 allocates nothing for them)", (long long)kNormPartsMin);
    // once per device, with that device current: the scratch's device address (the kernel's shared-memory opt-in and occupancy are
    // cached per device by norm_parts_cfg)
    static unsigned char *scratch_ptr[64] = {};
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    PCP_REQUIRE(dev >= 0 && dev < 64, PCP_ERR_UNSUPPORTED, "pcp_normalize_noalloc_f32: device ordinal %d", dev);
    if (!scratch_ptr[dev]) {
        void *p = nullptr;
        PCP_CUDA_OK(cudaGetSymbolAddress(&p, g_norm_noalloc_scratch));
        PCP_REQUIRE(p != nullptr, PCP_ERR_CUDA, "pcp_normalize_noalloc_f32: cudaGetSymbolAddress returned a null address");
        scratch_ptr[dev] = static_cast<unsigned char *>(p);
    }
    // the array keeps the previous call's contents: norm_parts_launch clears the ticket and this call's b records on the stream
    return norm_parts_launch(d_x, b, n, d_y, d_centroid, d_scale, scratch_ptr[dev], as_stream(stream));
}
