// sampling.cu — farthest-point sampling (north_star operator; no reference implementation, the
// contract is oracle/sampling.py:fps — SURVEY.md §8 a9).
//
// Design: FPS is a serial chain of S global arg-max steps, so the working set must never leave the
// chip.  One thread-block CLUSTER per cloud; every thread keeps PPT points (x,y,z) and their running
// min-distance in REGISTERS for the whole kernel (the cloud is read from HBM exactly once).  Per
// iteration: register-only distance update (3 sub, 3 mul, 2 add, min — no FMA, bit-exact with the
// oracle), REDUX warp arg-max, one __syncthreads, one 32-byte DSMEM record per CTA pushed to every
// peer, one cluster barrier.  Ties resolve to the lowest index (np.argmax).
#include <cooperative_groups.h>

#include <cstdlib>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace pcp {

constexpr int kFpsThreads = 512;
constexpr int kFpsWarps = kFpsThreads / 32;
constexpr int kFpsMaxCluster = 16;

struct __align__(16) FpsRec {
    float x, y, z, pad;       // one 16-byte read for the winner's coordinates
    unsigned long long key;   // (float bits of min-distance << 32) | (0xFFFFFFFF - point index)
    unsigned long long pad2;
};

__device__ __forceinline__ float sqdist_nofma(float px, float py, float pz, float cx, float cy, float cz) {
    const float dx = __fsub_rn(px, cx), dy = __fsub_rn(py, cy), dz = __fsub_rn(pz, cz);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}

// phase cycle totals of block 0 / thread 0 (development aid, read through pcp_fps_phase_cycles)
__device__ unsigned long long g_fps_phase[8];
#define FPS_STAMP(i) do { if (kProf && blockIdx.x == 0 && tid == 0) { const long long t_ = clock64(); ph[i] += t_ - tprev; tprev = t_; } } while (0)

template <int PPT, bool kProf>
__global__ void __launch_bounds__(kFpsThreads, 1)
fps_cluster_kernel(const float *__restrict__ x, int64_t n, int s, int start, int32_t *__restrict__ idx_out, int cs, int64_t ppc)
{
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
    float cx = __ldg(xc + 3 * (int64_t)start), cy = __ldg(xc + 3 * (int64_t)start + 1), cz = __ldg(xc + 3 * (int64_t)start + 2);
    if (rank == 0 && tid == 0) idx_out[cloud * s] = start;

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
        {
            // every warp reduces the 16 warp keys itself (no second barrier); the one thread that owns the block's winner
            // fetches its coordinates and pushes the record into every CTA of the cluster
            const unsigned long long key = lane < kFpsWarps ? warp_key[lane] : 0ull;
            const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
            const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
            const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
            const bool mine = mlo != 0u && cand == 0xFFFFFFFFu - mlo;
            if (mine || (mlo == 0u && tid == 0)) {
                FpsRec r;
                r.key = ((unsigned long long)mhi << 32) | mlo;
                r.x = r.y = r.z = r.pad = 0.f;
                if (mine) {
#pragma unroll
                    for (int k = 0; k < PPT; ++k)
                        if (k == bk) { r.x = px[k]; r.y = py[k]; r.z = pz[k]; }
                }
                if (cs == 1) slots[par][0] = r;
                else
                    for (int t = 0; t < cs; ++t) *cluster.map_shared_rank(&slots[par][rank], t) = r;
            }
        }
        FPS_STAMP(3);
        if (cs == 1) __syncthreads(); else cluster.sync();
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

template <int PPT, bool kProf = false>
static int launch_fps(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, int cs, int64_t ppc, cudaStream_t st)
{
    static bool attr_set = false;
    if (!attr_set) {
        PCP_CUDA_OK(cudaFuncSetAttribute(fps_cluster_kernel<PPT, kProf>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr_set = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((int64_t)b * cs));
    cfg.blockDim = dim3(kFpsThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, fps_cluster_kernel<PPT, kProf>, d_x, n, s, start, d_idx, cs, ppc));
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
extern "C" int pcp_fps_phase_cycles(unsigned long long out[6])
{
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpyFromSymbol(out, g_fps_phase, 6 * sizeof(unsigned long long)));
    return PCP_OK;
}
