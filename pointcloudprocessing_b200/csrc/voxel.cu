// voxel.cu — voxel-grid downsampling (north_star operator; no reference implementation, the
// contract is oracle/sampling.py:voxel_downsample — SURVEY.md §8 a10).
//
// Pipeline (all clouds in one launch each, everything asynchronous on the caller's stream):
//   1. per-cloud minimum corner (when no origin is given)             12 B/point read
//   2. key = floor((p - origin) / voxel) per axis -> 63-bit key, payload = point index
//   3. stable LSD radix sort of (key, index) per cloud, 8 passes x 8 bits, warp match-any ranking
//   4. run heads -> voxel ids (per-cloud scan) -> one thread per voxel sums its run in double
// Double accumulation makes the centroid independent of summation order (bit-stable results).
//
// Fast path (taken when every cloud's occupied bounding grid has at most 32 n cells — LiDAR scans at 0.5-1 m voxels are
// two orders of magnitude below that): no sort at all.  The voxel's rank in (kx, ky, kz) order IS the number of occupied
// cells before it in the dense grid, so
//   1. per-cloud min / max corner                         12 B/point read
//   2. occupancy bitmap of the bounding grid (atomicOr)    12 B/point read
//   3. per-cloud popcount scan of the bitmap words -> rank of every occupied cell, voxel count
//   4. per point: rank lookup, fp64 atomic adds of x, y, z and the count into the voxel's accumulators (12 B/point read)
//   5. centroids = sum / count, zero padding               16 B/output row written
// Same keys, counts and centroids as the sort pipeline (exact integer work; the fp64 sums of fp32 values are exact at
// these magnitudes, hence order independent).  The dispatch reads one flag back from the device (a ~10 us
// synchronisation on the caller's stream); the sort pipeline below remains the general path.
#include "common.cuh"

namespace pcp {

constexpr int kSortThreads = 256;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;   // 4096 elements per block
constexpr int kKeyBias = 1 << 20;

__device__ __forceinline__ unsigned enc_f32(float f) {
    const unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float dec_f32(unsigned e) {
    return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e);
}

__global__ void __launch_bounds__(256)
voxel_min_kernel(const float *__restrict__ x, int64_t n, unsigned *__restrict__ origin_enc)
{
    const int64_t cloud = blockIdx.y;
    const float *xc = x + cloud * n * 3;
    float m[3] = {__int_as_float(0x7f800000), __int_as_float(0x7f800000), __int_as_float(0x7f800000)};
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        m[0] = fminf(m[0], xc[3 * i]); m[1] = fminf(m[1], xc[3 * i + 1]); m[2] = fminf(m[2], xc[3 * i + 2]);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m[a] = fminf(m[a], __shfl_xor_sync(0xffffffffu, m[a], o));
        if ((threadIdx.x & 31) == 0) atomicMin(origin_enc + cloud * 3 + a, enc_f32(m[a]));
    }
}

__global__ void __launch_bounds__(256)
voxel_key_kernel(const float *__restrict__ x, int64_t n, const unsigned *__restrict__ origin_enc, float ox, float oy, float oz,
                 int use_enc, float voxel, unsigned long long *__restrict__ keys, unsigned *__restrict__ vals,
                 int *__restrict__ bad)
{
    const int64_t cloud = blockIdx.y;
    float o[3] = {ox, oy, oz};
    if (use_enc) { o[0] = dec_f32(origin_enc[cloud * 3]); o[1] = dec_f32(origin_enc[cloud * 3 + 1]); o[2] = dec_f32(origin_enc[cloud * 3 + 2]); }
    const float *xc = x + cloud * n * 3;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        unsigned long long key = 0;
        bool oob = false;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float q = __fdiv_rn(__fsub_rn(xc[3 * i + a], o[a]), voxel);
            const float fl = floorf(q);
            oob |= !(fl > -(float)kKeyBias && fl < (float)kKeyBias);
            const long long k = (long long)fl + kKeyBias;
            key = (key << 21) | (unsigned long long)(k & 0x1FFFFF);
        }
        if (oob) bad[cloud] = 1;
        keys[cloud * n + i] = key;
        vals[cloud * n + i] = (unsigned)i;
    }
}

// ---- radix sort pass: histogram -> per-cloud scan -> stable scatter ----
__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const unsigned long long *__restrict__ keys, int64_t n, int shift, unsigned *__restrict__ hist, int nb)
{
    __shared__ unsigned h[256];
    const int64_t cloud = blockIdx.y;
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
    for (int j = 0; j < kSortItems; ++j) {
        const int64_t i = base + (int64_t)j * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[cloud * n + i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(cloud * 256 + threadIdx.x) * nb + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan over (digit-major, block-minor) counts of one cloud; in place
__global__ void __launch_bounds__(1024)
sort_scan_kernel(unsigned *__restrict__ hist, int nb)
{
    __shared__ unsigned part[1024];
    unsigned *h = hist + (int64_t)blockIdx.x * 256 * nb;
    const int total = 256 * nb;
    const int per = (total + 1023) / 1024;
    const int beg = threadIdx.x * per, end = min(total, beg + per);
    unsigned s = 0;
    for (int i = beg; i < end; ++i) s += h[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        unsigned v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = part[threadIdx.x] - s;
    for (int i = beg; i < end; ++i) { const unsigned c = h[i]; h[i] = run; run += c; }
}

__global__ void __launch_bounds__(kSortThreads)
sort_scatter_kernel(const unsigned long long *__restrict__ keys_in, const unsigned *__restrict__ vals_in,
                    unsigned long long *__restrict__ keys_out, unsigned *__restrict__ vals_out, int64_t n, int shift,
                    const unsigned *__restrict__ hist, int nb)
{
    constexpr int kWarps = kSortThreads / 32;
    __shared__ unsigned wcount[kWarps][256];
    const int64_t cloud = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < kWarps * 256; i += kSortThreads) (&wcount[0][0])[i] = 0;
    __syncthreads();
    // warp w owns the contiguous range [base + w*512, base + (w+1)*512); round r covers 32 consecutive elements
    const int64_t wbase = (int64_t)blockIdx.x * kSortTile + (int64_t)warp * (32 * kSortItems);
    unsigned long long k[kSortItems];
    unsigned v[kSortItems], rank[kSortItems];
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        const bool ok = i < n;
        k[r] = ok ? keys_in[cloud * n + i] : ~0ull;
        v[r] = ok ? vals_in[cloud * n + i] : 0u;
        const unsigned d = ok ? ((unsigned)(k[r] >> shift) & 255u) : 256u;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned before = __popc(peers & ((1u << lane) - 1u));
        unsigned basec = 0;
        if (ok) basec = wcount[warp][d];
        __syncwarp();
        rank[r] = basec + before;
        if (ok && before == 0) wcount[warp][d] = basec + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // exclusive prefix over warps for digit tid, plus the block's global offset for that digit
    {
        unsigned run = hist[(cloud * 256 + tid) * nb + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kWarps; ++w) { const unsigned c = wcount[w][tid]; wcount[w][tid] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortItems; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        if (i < n) {
            const unsigned d = (unsigned)(k[r] >> shift) & 255u;
            const int64_t pos = cloud * n + wcount[warp][d] + rank[r];
            keys_out[pos] = k[r];
            vals_out[pos] = v[r];
        }
    }
}

// run heads -> voxel ids.  One block per cloud: flag, block-wide scan, starts[vid] = position.
__global__ void __launch_bounds__(1024)
voxel_heads_kernel(const unsigned long long *__restrict__ keys, int64_t n, unsigned *__restrict__ starts,
                   int32_t *__restrict__ nvox, const int *__restrict__ bad)
{
    __shared__ unsigned part[1024];
    const int64_t cloud = blockIdx.x;
    const unsigned long long *kc = keys + cloud * n;
    const int64_t per = (n + 1023) / 1024;
    const int64_t beg = threadIdx.x * per, end = min(n, beg + per);
    unsigned s = 0;
    for (int64_t i = beg; i < end; ++i) s += (i == 0 || kc[i] != kc[i - 1]) ? 1u : 0u;
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        unsigned v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned vid = part[threadIdx.x] - s;
    for (int64_t i = beg; i < end; ++i)
        if (i == 0 || kc[i] != kc[i - 1]) starts[cloud * (n + 1) + vid++] = (unsigned)i;
    if (threadIdx.x == 1023) {
        const unsigned total = part[1023];
        starts[cloud * (n + 1) + total] = (unsigned)n;
        nvox[cloud] = bad[cloud] ? -1 : (int32_t)total;
    }
}

__global__ void __launch_bounds__(256)
voxel_reduce_kernel(const float *__restrict__ x, int64_t n, const unsigned *__restrict__ vals, const unsigned *__restrict__ starts,
                    const int32_t *__restrict__ nvox, float *__restrict__ out_xyz, int32_t *__restrict__ out_count)
{
    const int64_t cloud = blockIdx.y;
    const int nv = max(nvox[cloud], 0);
    const float *xc = x + cloud * n * 3;
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
        float cx = 0.f, cy = 0.f, cz = 0.f;
        int cnt = 0;
        if (v < nv) {
            const unsigned s0 = starts[cloud * (n + 1) + v], s1 = starts[cloud * (n + 1) + v + 1];
            double sx = 0, sy = 0, sz = 0;
            for (unsigned i = s0; i < s1; ++i) {
                const unsigned p = vals[cloud * n + i];
                sx += (double)xc[3 * (int64_t)p]; sy += (double)xc[3 * (int64_t)p + 1]; sz += (double)xc[3 * (int64_t)p + 2];
            }
            cnt = (int)(s1 - s0);
            cx = (float)(sx / (double)cnt); cy = (float)(sy / (double)cnt); cz = (float)(sz / (double)cnt);
        }
        float *o = out_xyz + (cloud * n + v) * 3;
        o[0] = cx; o[1] = cy; o[2] = cz;
        out_count[cloud * n + v] = cnt;
    }
}

// ---- fast path: occupancy bitmap + popcount ranks ----
struct VoxelGrid {                 // per cloud
    float o[3];                    // origin
    int kmin[3];                   // smallest key per axis
    int ny, nz;                    // grid extents of the y and z axes
    long long cells;               // nx * ny * nz
};

__global__ void __launch_bounds__(256)
voxel_minmax_kernel(const float *__restrict__ x, int64_t n, unsigned *__restrict__ min_enc, unsigned *__restrict__ max_enc, int vec_ok)
{
    const int64_t cloud = blockIdx.y;
    const float *xc = x + cloud * n * 3;
    const float inf = __int_as_float(0x7f800000);
    float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
    if (vec_ok) {
        // lane-contiguous float4 loads; load slot u's component j is axis (u T + gt + j) % 3 on every iteration (stride 3 T)
        const float4 *x4 = reinterpret_cast<const float4 *>(xc);
        const int64_t nvec = (n * 3) >> 2, T = (int64_t)gridDim.x * 256, gt = (int64_t)blockIdx.x * 256 + threadIdx.x;
        float l[3][4], h[3][4];
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) { l[u][j] = inf; h[u][j] = -inf; }
        for (int64_t v0 = gt; v0 < nvec; v0 += 3 * T) {
            float4 q[3];
#pragma unroll
            for (int u = 0; u < 3; ++u) q[u] = (v0 + u * T < nvec) ? __ldg(x4 + v0 + u * T) : make_float4(inf, inf, inf, inf);
#pragma unroll
            for (int u = 0; u < 3; ++u) {
                const bool ok = v0 + u * T < nvec;
                l[u][0] = fminf(l[u][0], q[u].x); l[u][1] = fminf(l[u][1], q[u].y); l[u][2] = fminf(l[u][2], q[u].z); l[u][3] = fminf(l[u][3], q[u].w);
                if (ok) { h[u][0] = fmaxf(h[u][0], q[u].x); h[u][1] = fmaxf(h[u][1], q[u].y); h[u][2] = fmaxf(h[u][2], q[u].z); h[u][3] = fmaxf(h[u][3], q[u].w); }
            }
        }
#pragma unroll
        for (int u = 0; u < 3; ++u)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int axis = (int)((u * T + gt + j) % 3);
#pragma unroll
                for (int a = 0; a < 3; ++a)
                    if (axis == a) { lo[a] = fminf(lo[a], l[u][j]); hi[a] = fmaxf(hi[a], h[u][j]); }
            }
    } else {
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
            for (int a = 0; a < 3; ++a) { const float v = xc[3 * i + a]; lo[a] = fminf(lo[a], v); hi[a] = fmaxf(hi[a], v); }
        }
    }
    __shared__ float red[2][8][3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5][a] = lo[a]; red[1][threadIdx.x >> 5][a] = hi[a]; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        float l = inf, h = -inf;
        for (int w = 0; w < 8; ++w) { l = fminf(l, red[0][w][threadIdx.x]); h = fmaxf(h, red[1][w][threadIdx.x]); }
        atomicMin(min_enc + cloud * 3 + threadIdx.x, enc_f32(l));
        atomicMax(max_enc + cloud * 3 + threadIdx.x, enc_f32(h));
    }
}

__device__ __forceinline__ float voxel_keyf(float v, float o, float voxel) { return floorf(__fdiv_rn(__fsub_rn(v, o), voxel)); }

// one thread per cloud: grid of the occupied bounding box; *need_sort = 1 when a cloud does not fit the bitmap
__global__ void voxel_grid_kernel(const unsigned *__restrict__ min_enc, const unsigned *__restrict__ max_enc, int b, float ox, float oy, float oz,
                                  int use_min, float voxel, long long cap_cells, VoxelGrid *__restrict__ grid, int *__restrict__ need_sort)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= b) return;
    VoxelGrid g;
    const float og[3] = {ox, oy, oz};
    long long cells = 1;
    bool ok = true;
    int dim[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = dec_f32(min_enc[c * 3 + a]), hi = dec_f32(max_enc[c * 3 + a]);
        g.o[a] = use_min ? lo : og[a];
        const float k0 = voxel_keyf(lo, g.o[a], voxel), k1 = voxel_keyf(hi, g.o[a], voxel);
        ok = ok && (k0 > -(float)kKeyBias) && (k1 < (float)kKeyBias) && (k1 >= k0);
        g.kmin[a] = ok ? (int)k0 : 0;
        dim[a] = ok ? (int)k1 - (int)k0 + 1 : 1;
        cells = (cells <= cap_cells) ? cells * dim[a] : cells;
    }
    g.ny = dim[1]; g.nz = dim[2]; g.cells = cells;
    grid[c] = g;
    if (!ok || cells > cap_cells) *need_sort = 1;
}

__device__ __forceinline__ long long voxel_cell(const float *p, const VoxelGrid &g, float voxel, bool &finite)
{
    const float kx = voxel_keyf(p[0], g.o[0], voxel), ky = voxel_keyf(p[1], g.o[1], voxel), kz = voxel_keyf(p[2], g.o[2], voxel);
    finite = (kx == kx) && (ky == ky) && (kz == kz);         // NaN coordinates (min / max ignore them)
    return ((long long)((int)kx - g.kmin[0]) * g.ny + ((int)ky - g.kmin[1])) * g.nz + ((int)kz - g.kmin[2]);
}

// The fast-path kernels are enqueued right behind voxel_grid_kernel and return at once when it raised *need_sort: the host
// reads the flag only after everything is queued, so the decision costs no bubble on the device.

// zero the bitmap words the cloud's grid actually uses (the capacity is 32 n cells per cloud)
__global__ void __launch_bounds__(256)
voxel_clear_kernel(const VoxelGrid *__restrict__ grid, long long cap_cells, unsigned *__restrict__ bitmap, const int *__restrict__ need_sort)
{
    if (*need_sort) return;
    const int64_t cloud = blockIdx.y;
    const long long vec = (((grid[cloud].cells + 31) >> 5) + 3) >> 2;          // capacity is a multiple of 32 words
    uint4 *bm = reinterpret_cast<uint4 *>(bitmap + cloud * (cap_cells >> 5));
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < vec; i += (long long)gridDim.x * blockDim.x)
        bm[i] = make_uint4(0u, 0u, 0u, 0u);
}

// four points per thread from three 16-byte loads when the cloud is 16-byte aligned (n % 4 == 0)
template <typename F>
__device__ __forceinline__ void voxel_for_points(const float *__restrict__ xc, int64_t n, int vec_ok, F &&f)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t ngroups = vec_ok ? (n >> 2) : 0;
    const float4 *x4 = reinterpret_cast<const float4 *>(xc);
    for (int64_t g = t0; g < ngroups; g += stride) {
        const float4 a = __ldg(x4 + 3 * g), b = __ldg(x4 + 3 * g + 1), c = __ldg(x4 + 3 * g + 2);
        const float p0[3] = {a.x, a.y, a.z}, p1[3] = {a.w, b.x, b.y}, p2[3] = {b.z, b.w, c.x}, p3[3] = {c.y, c.z, c.w};
        f(4 * g, p0); f(4 * g + 1, p1); f(4 * g + 2, p2); f(4 * g + 3, p3);
    }
    for (int64_t i = (ngroups << 2) + t0; i < n; i += stride) {
        const float p[3] = {xc[3 * i], xc[3 * i + 1], xc[3 * i + 2]};
        f(i, p);
    }
}

__global__ void __launch_bounds__(256)
voxel_mark_kernel(const float *__restrict__ x, int64_t n, const VoxelGrid *__restrict__ grid, float voxel, long long cap_cells,
                  unsigned *__restrict__ bitmap, int *__restrict__ bad, const int *__restrict__ need_sort, int vec_ok)
{
    if (*need_sort) return;
    const int64_t cloud = blockIdx.y;
    const VoxelGrid g = grid[cloud];
    unsigned *bm = bitmap + cloud * (cap_cells >> 5);
    // ~8 points share a voxel and ~50 a word: look before the read-modify-write (L2 read, the line may be stale in L1).  The
    // four look-ups of a thread's points are issued together — behind each other's atomics they would be four serial L2
    // round trips (long-scoreboard stalls were 27 of 40 issue slots, profiles/r01_preproc_ncu_v10.txt).
    const float *xc = x + cloud * n * 3;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t ngroups = vec_ok ? (n >> 2) : 0;
    const float4 *x4 = reinterpret_cast<const float4 *>(xc);
    for (int64_t gq = t0; gq < ngroups; gq += stride) {
        const float4 a = __ldg(x4 + 3 * gq), b4 = __ldg(x4 + 3 * gq + 1), c = __ldg(x4 + 3 * gq + 2);
        const float p[4][3] = {{a.x, a.y, a.z}, {a.w, b4.x, b4.y}, {b4.z, b4.w, c.x}, {c.y, c.z, c.w}};
        long long cell[4];
        bool fin[4];
        unsigned w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) cell[j] = voxel_cell(p[j], g, voxel, fin[j]);
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = fin[j] ? __ldcg(bm + (cell[j] >> 5)) : 0xFFFFFFFFu;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (!fin[j]) { bad[cloud] = 1; continue; }
            const unsigned bit = 1u << (cell[j] & 31);
            if (!(w[j] & bit)) atomicOr(bm + (cell[j] >> 5), bit);
        }
    }
    for (int64_t i = (ngroups << 2) + t0; i < n; i += stride) {
        const float p[3] = {xc[3 * i], xc[3 * i + 1], xc[3 * i + 2]};
        bool finite;
        const long long cell = voxel_cell(p, g, voxel, finite);
        if (!finite) { bad[cloud] = 1; continue; }
        const unsigned bit = 1u << (cell & 31);
        if (!(__ldcg(bm + (cell >> 5)) & bit)) atomicOr(bm + (cell >> 5), bit);
    }
}

// one CTA per cloud: exclusive popcount scan over the bitmap words -> rank of the first cell of every word; voxel count;
// accumulators of the occupied voxels zeroed.  A thread owns eight consecutive words (two 16-byte loads), so one block
// scan covers 8192 words.
__global__ void __launch_bounds__(1024)
voxel_rank_kernel(const unsigned *__restrict__ bitmap, const VoxelGrid *__restrict__ grid, long long cap_cells, int64_t n,
                  unsigned *__restrict__ word_rank, int32_t *__restrict__ nvox, const int *__restrict__ bad, double *__restrict__ sums,
                  int *__restrict__ cnt, const int *__restrict__ need_sort)
{
    if (*need_sort) return;
    __shared__ unsigned wtot[32];
    __shared__ unsigned carry_s;
    const int64_t cloud = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long vec = (((grid[cloud].cells + 31) >> 5) + 3) >> 2;          // uint4 count
    const uint4 *bm = reinterpret_cast<const uint4 *>(bitmap + cloud * (cap_cells >> 5));
    uint4 *wr = reinterpret_cast<uint4 *>(word_rank + cloud * (cap_cells >> 5));
    unsigned carry = 0;
    for (long long base = 0; base < vec; base += 2048) {
        const long long i0 = base + 2 * tid;
        uint4 q0 = make_uint4(0u, 0u, 0u, 0u), q1 = q0;
        if (i0 < vec) q0 = bm[i0];
        if (i0 + 1 < vec) q1 = bm[i0 + 1];
        const unsigned c[8] = {(unsigned)__popc(q0.x), (unsigned)__popc(q0.y), (unsigned)__popc(q0.z), (unsigned)__popc(q0.w),
                               (unsigned)__popc(q1.x), (unsigned)__popc(q1.y), (unsigned)__popc(q1.z), (unsigned)__popc(q1.w)};
        unsigned tsum = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) tsum += c[k];
        unsigned incl = tsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        if (lane == 31) wtot[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            unsigned w = wtot[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += v; }
            wtot[lane] = wi - w;
            if (lane == 31) carry_s = wi;
        }
        __syncthreads();
        unsigned r = carry + wtot[warp] + incl - tsum;
        carry += carry_s;
        unsigned e[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { e[k] = r; r += c[k]; }
        if (i0 < vec) wr[i0] = make_uint4(e[0], e[1], e[2], e[3]);
        if (i0 + 1 < vec) wr[i0 + 1] = make_uint4(e[4], e[5], e[6], e[7]);
        __syncthreads();
    }
    const unsigned total = carry;
    if (tid == 0) nvox[cloud] = bad[cloud] ? -1 : (int32_t)total;
}

// accumulators of the occupied voxels zeroed by the whole GPU (one CTA per cloud took as long over it as the scan itself)
__global__ void __launch_bounds__(256)
voxel_zero_kernel(int64_t n, const int32_t *__restrict__ nvox, double *__restrict__ sums, const int *__restrict__ need_sort)
{
    if (*need_sort) return;
    const int64_t cloud = blockIdx.y;
    const int64_t nv = max(nvox[cloud], 0);
    double4 *sp = reinterpret_cast<double4 *>(sums) + cloud * n;                // {sum x, sum y, sum z, count}: one 32-byte sector
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nv; v += (int64_t)gridDim.x * blockDim.x)
        sp[v] = make_double4(0.0, 0.0, 0.0, 0.0);
}

// A voxel's accumulators {sum x, sum y, sum z, count} fill one 32-byte sector, and the four reductions of a point are
// issued by four neighbouring lanes of ONE instruction: the L2 sees a quarter of the sector operations that per-thread
// atomicAdds (one instruction per component, 32 sectors each) generate — the kernel was bound by exactly that
// (lts__throughput 66 %, lg_throttle + long_scoreboard 75 of 80 stalled slots, profiles/r01_preproc_ncu_v10.txt).
__global__ void __launch_bounds__(256)
voxel_accum_kernel(const float *__restrict__ x, int64_t n, const VoxelGrid *__restrict__ grid, float voxel, long long cap_cells,
                   const unsigned *__restrict__ bitmap, const unsigned *__restrict__ word_rank, double *__restrict__ sums, int *__restrict__ cnt,
                   const int *__restrict__ need_sort, int vec_ok)
{
    if (*need_sort) return;
    (void)cnt;
    __shared__ float4 tile[8][128];                         // per warp: {x, y, z, voxel rank bits} of 128 points
    const int64_t cloud = blockIdx.y;
    const VoxelGrid g = grid[cloud];
    const unsigned *bm = bitmap + cloud * (cap_cells >> 5), *wr = word_rank + cloud * (cap_cells >> 5);
    const float *xc = x + cloud * n * 3;
    double *acc = sums + cloud * n * 4;
    const int lane = threadIdx.x & 31;
    float4 *tw = tile[threadIdx.x >> 5];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    const int64_t ngroups = vec_ok ? (n >> 2) : 0, nfull = (ngroups + 31) & ~(int64_t)31;      // whole warps stay in the loop together
    const float4 *x4 = reinterpret_cast<const float4 *>(xc);
    auto rank_of = [&](const float *p) -> unsigned {
        bool finite;
        const long long cell = voxel_cell(p, g, voxel, finite);
        if (!finite) return 0xFFFFFFFFu;
        const unsigned w = __ldg(bm + (cell >> 5));
        return __ldg(wr + (cell >> 5)) + __popc(w & ((1u << (cell & 31)) - 1u));
    };
    for (int64_t gq = t0; gq < nfull; gq += stride) {
        float p[4][3] = {};
        unsigned v[4] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu};
        if (gq < ngroups) {
            const float4 a = __ldg(x4 + 3 * gq), b4 = __ldg(x4 + 3 * gq + 1), c = __ldg(x4 + 3 * gq + 2);
            p[0][0] = a.x; p[0][1] = a.y; p[0][2] = a.z; p[1][0] = a.w; p[1][1] = b4.x; p[1][2] = b4.y;
            p[2][0] = b4.z; p[2][1] = b4.w; p[2][2] = c.x; p[3][0] = c.y; p[3][1] = c.z; p[3][2] = c.w;
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = rank_of(p[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) tw[j * 32 + lane] = make_float4(p[j][0], p[j][1], p[j][2], __uint_as_float(v[j]));
        __syncwarp();
        const int comp = lane & 3;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            const float4 e = tw[q * 8 + (lane >> 2)];
            const unsigned vv = __float_as_uint(e.w);
            const double val = comp == 0 ? (double)e.x : comp == 1 ? (double)e.y : comp == 2 ? (double)e.z : 1.0;
            if (vv != 0xFFFFFFFFu) atomicAdd(acc + (int64_t)vv * 4 + comp, val);
        }
        __syncwarp();
    }
    for (int64_t i = (ngroups << 2) + t0; i < n; i += stride) {
        const float p[3] = {xc[3 * i], xc[3 * i + 1], xc[3 * i + 2]};
        const unsigned vv = rank_of(p);
        if (vv == 0xFFFFFFFFu) continue;
        atomicAdd(acc + (int64_t)vv * 4, (double)p[0]); atomicAdd(acc + (int64_t)vv * 4 + 1, (double)p[1]);
        atomicAdd(acc + (int64_t)vv * 4 + 2, (double)p[2]); atomicAdd(acc + (int64_t)vv * 4 + 3, 1.0);
    }
}

__global__ void __launch_bounds__(256)
voxel_finalize_kernel(int64_t n, const int32_t *__restrict__ nvox, const double *__restrict__ sums, const int *__restrict__ cnt,
                      float *__restrict__ out_xyz, int32_t *__restrict__ out_count, const int *__restrict__ need_sort)
{
    if (*need_sort) return;
    const int64_t cloud = blockIdx.y;
    const int nv = max(nvox[cloud], 0);
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
        float cx = 0.f, cy = 0.f, cz = 0.f;
        int c = 0;
        if (v < nv) {
            const double4 sp = reinterpret_cast<const double4 *>(sums)[cloud * n + v];
            c = (int)sp.w;                                   // counts are exact in a double
            cx = (float)(sp.x / sp.w); cy = (float)(sp.y / sp.w); cz = (float)(sp.z / sp.w);
        }
        float *o = out_xyz + (cloud * n + v) * 3;
        o[0] = cx; o[1] = cy; o[2] = cz;
        out_count[cloud * n + v] = c;
    }
}

// compact form of the finalize pass: rows [0, min(nvox, capacity)) only — no zero padding (at 64 x 100 000 points the padded
// b x n rows were 102 MB of writes for 12 MB of voxels); a batch the bitmap path does not cover reports nvox = -2
__global__ void __launch_bounds__(256)
voxel_finalize_compact_kernel(int64_t n, int64_t capacity, int32_t *__restrict__ nvox, const double *__restrict__ sums, float *__restrict__ out_xyz,
                              int32_t *__restrict__ out_count, const int *__restrict__ need_sort)
{
    const int64_t cloud = blockIdx.y;
    if (*need_sort) { if (blockIdx.x == 0 && threadIdx.x == 0) nvox[cloud] = -2; return; }
    const int64_t nv = min((int64_t)max(nvox[cloud], 0), capacity);
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < nv; v += (int64_t)gridDim.x * blockDim.x) {
        const double4 sp = reinterpret_cast<const double4 *>(sums)[cloud * n + v];
        float *o = out_xyz + (cloud * capacity + v) * 3;
        o[0] = (float)(sp.x / sp.w); o[1] = (float)(sp.y / sp.w); o[2] = (float)(sp.z / sp.w);
        out_count[cloud * capacity + v] = (int)sp.w;                 // counts are exact in a double
    }
}

struct VoxelWs {
    unsigned long long *keys_a, *keys_b;
    unsigned *vals_a, *vals_b, *hist, *starts, *origin_enc;
    int *bad;
    // fast path
    unsigned *max_enc, *bitmap, *word_rank;
    VoxelGrid *grid;
    double *sums;
    int *cnt, *need_sort;
    long long cap_cells;
    size_t bytes;
};

static VoxelWs carve(void *base, int b, int64_t n)
{
    VoxelWs w;
    const size_t bn = (size_t)b * (size_t)n;
    const int nb = (int)((n + kSortTile - 1) / kSortTile);
    char *p = reinterpret_cast<char *>(base);
    auto take = [&](size_t bytes) { char *q = p; p += (bytes + 255) & ~(size_t)255; return q; };
    w.keys_a = reinterpret_cast<unsigned long long *>(take(bn * 8));
    w.keys_b = reinterpret_cast<unsigned long long *>(take(bn * 8));
    w.vals_a = reinterpret_cast<unsigned *>(take(bn * 4));
    w.vals_b = reinterpret_cast<unsigned *>(take(bn * 4));
    w.hist = reinterpret_cast<unsigned *>(take((size_t)b * 256 * nb * 4));
    w.starts = reinterpret_cast<unsigned *>(take((size_t)b * (n + 1) * 4));
    w.origin_enc = reinterpret_cast<unsigned *>(take((size_t)b * 3 * 4));
    w.bad = reinterpret_cast<int *>(take((size_t)b * 4));
    w.cap_cells = ((32 * (long long)n + 1023) / 1024) * 1024;
    w.max_enc = reinterpret_cast<unsigned *>(take((size_t)b * 3 * 4));
    w.bitmap = reinterpret_cast<unsigned *>(take((size_t)b * (size_t)(w.cap_cells >> 3)));
    w.word_rank = reinterpret_cast<unsigned *>(take((size_t)b * (size_t)(w.cap_cells >> 3)));
    w.grid = reinterpret_cast<VoxelGrid *>(take((size_t)b * sizeof(VoxelGrid)));
    w.sums = reinterpret_cast<double *>(take(bn * 32));
    w.cnt = reinterpret_cast<int *>(take(bn * 4));
    w.need_sort = reinterpret_cast<int *>(take(4));
    w.bytes = (size_t)(p - reinterpret_cast<char *>(base));
    return w;
}

// grid.x of a (gx, b) grid-stride launch that is resident in ONE wave: a grid of sm_count * 8 CTAs on a kernel that fits five
// or seven per SM runs a second, mostly empty wave (1152 CTAs on 740-1036 slots at 64 clouds)
template <typename K>
static int wave_gx(K kernel, int b, int64_t items_per_cloud, int *cache)
{
    if (*cache == 0) {
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 256, 0) != cudaSuccess || occ < 1) occ = 4;
        *cache = occ;
    }
    const int64_t want = std::max<int64_t>(1, (int64_t)sm_count() * *cache / b);
    return (int)std::max<int64_t>(1, std::min<int64_t>((items_per_cloud + 255) / 256, want));
}

}  // namespace pcp

using namespace pcp;

extern "C" int pcp_voxel_workspace_bytes(int b, int64_t n, size_t *bytes)
{
    PCP_REQUIRE(bytes && b >= 0 && n >= 0, PCP_ERR_INVALID, "pcp_voxel_workspace_bytes: bad arguments");
    *bytes = carve(nullptr, b > 0 ? b : 1, n > 0 ? n : 1).bytes;
    return PCP_OK;
}

extern "C" int pcp_voxel_f32(const float *d_x, int b, int64_t n, const float *h_origin, float voxel, float *d_out_xyz,
                             int32_t *d_out_count, int32_t *d_nvox, void *d_workspace, size_t workspace_bytes, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_out_xyz && d_out_count && d_nvox && d_workspace && b >= 0 && n >= 1, PCP_ERR_INVALID, "pcp_voxel_f32: bad arguments");
    PCP_REQUIRE(voxel > 0.f, PCP_ERR_INVALID, "pcp_voxel_f32: voxel size must be > 0");
    PCP_REQUIRE(n < 0xFFFFFFFFll && b <= 65535, PCP_ERR_UNSUPPORTED, "pcp_voxel_f32: n or b too large");
    if (b == 0) return PCP_OK;
    VoxelWs w = carve(d_workspace, b, n);
    PCP_REQUIRE(w.bytes <= workspace_bytes, PCP_ERR_INVALID, "pcp_voxel_f32: workspace too small (%zu < %zu)", workspace_bytes, w.bytes);
    cudaStream_t st = as_stream(stream);
    const int nb = (int)((n + kSortTile - 1) / kSortTile);
    int gx = (int)std::min<int64_t>((n + 255) / 256, std::max(1, sm_count() * 8 / b));
    PCP_CUDA_OK(cudaMemsetAsync(w.bad, 0, (size_t)b * 4, st));
    // ---- fast path: bitmap ranks (see the header comment) ----
    {
        const float ox = h_origin ? h_origin[0] : 0.f, oy = h_origin ? h_origin[1] : 0.f, oz = h_origin ? h_origin[2] : 0.f;
        PCP_CUDA_OK(cudaMemsetAsync(w.origin_enc, 0xFF, (size_t)b * 12, st));
        PCP_CUDA_OK(cudaMemsetAsync(w.max_enc, 0, (size_t)b * 12, st));
        PCP_CUDA_OK(cudaMemsetAsync(w.need_sort, 0, 4, st));
        const int vec_ok = ((reinterpret_cast<uintptr_t>(d_x) & 15u) == 0) && (n % 4 == 0);
        static int occ_minmax = 0, occ_clear = 0, occ_mark = 0, occ_accum = 0, occ_fin = 0, occ_zero = 0;
        const int64_t groups = vec_ok ? (n + 3) / 4 : n;
        voxel_minmax_kernel<<<dim3(wave_gx(voxel_minmax_kernel, b, vec_ok ? (n * 3 / 4 + 2) / 3 : n, &occ_minmax), b), 256, 0, st>>>(d_x, n, w.origin_enc, w.max_enc, vec_ok);
        PCP_CHECK_LAUNCH();
        voxel_grid_kernel<<<(b + 127) / 128, 128, 0, st>>>(w.origin_enc, w.max_enc, b, ox, oy, oz, h_origin ? 0 : 1, voxel, w.cap_cells, w.grid,
                                                          w.need_sort);
        PCP_CHECK_LAUNCH();
        voxel_clear_kernel<<<dim3(wave_gx(voxel_clear_kernel, b, (w.cap_cells >> 7) + 1, &occ_clear), b), 256, 0, st>>>(w.grid, w.cap_cells, w.bitmap, w.need_sort);
        PCP_CHECK_LAUNCH();
        voxel_mark_kernel<<<dim3(wave_gx(voxel_mark_kernel, b, groups, &occ_mark), b), 256, 0, st>>>(d_x, n, w.grid, voxel, w.cap_cells, w.bitmap, w.bad, w.need_sort, vec_ok);
        PCP_CHECK_LAUNCH();
        voxel_rank_kernel<<<b, 1024, 0, st>>>(w.bitmap, w.grid, w.cap_cells, n, w.word_rank, d_nvox, w.bad, w.sums, w.cnt, w.need_sort);
        PCP_CHECK_LAUNCH();
        voxel_zero_kernel<<<dim3(wave_gx(voxel_zero_kernel, b, n, &occ_zero), b), 256, 0, st>>>(n, d_nvox, w.sums, w.need_sort);
        PCP_CHECK_LAUNCH();
        voxel_accum_kernel<<<dim3(wave_gx(voxel_accum_kernel, b, groups, &occ_accum), b), 256, 0, st>>>(d_x, n, w.grid, voxel, w.cap_cells, w.bitmap, w.word_rank, w.sums, w.cnt, w.need_sort, vec_ok);
        PCP_CHECK_LAUNCH();
        voxel_finalize_kernel<<<dim3(wave_gx(voxel_finalize_kernel, b, n, &occ_fin), b), 256, 0, st>>>(n, d_nvox, w.sums, w.cnt, d_out_xyz, d_out_count, w.need_sort);
        PCP_CHECK_LAUNCH();
        // the flag is read once everything is queued: the wait overlaps the kernels instead of stalling the stream mid-way
        int need_sort = 0;
        PCP_CUDA_OK(cudaMemcpyAsync(&need_sort, w.need_sort, 4, cudaMemcpyDeviceToHost, st));
        PCP_CUDA_OK(cudaStreamSynchronize(st));
        if (!need_sort) return PCP_OK;
    }
    // ---- general path: sort ----
    if (!h_origin) {
        PCP_CUDA_OK(cudaMemsetAsync(w.origin_enc, 0xFF, (size_t)b * 12, st));
        voxel_min_kernel<<<dim3(gx, b), 256, 0, st>>>(d_x, n, w.origin_enc);
        PCP_CHECK_LAUNCH();
    }
    voxel_key_kernel<<<dim3(gx, b), 256, 0, st>>>(d_x, n, w.origin_enc, h_origin ? h_origin[0] : 0.f, h_origin ? h_origin[1] : 0.f,
                                                  h_origin ? h_origin[2] : 0.f, h_origin ? 0 : 1, voxel, w.keys_a, w.vals_a, w.bad);
    PCP_CHECK_LAUNCH();
    unsigned long long *ka = w.keys_a, *kb = w.keys_b;
    unsigned *va = w.vals_a, *vb = w.vals_b;
    for (int shift = 0; shift < 64; shift += 8) {
        sort_hist_kernel<<<dim3(nb, b), kSortThreads, 0, st>>>(ka, n, shift, w.hist, nb);
        PCP_CHECK_LAUNCH();
        sort_scan_kernel<<<b, 1024, 0, st>>>(w.hist, nb);
        PCP_CHECK_LAUNCH();
        sort_scatter_kernel<<<dim3(nb, b), kSortThreads, 0, st>>>(ka, va, kb, vb, n, shift, w.hist, nb);
        PCP_CHECK_LAUNCH();
        std::swap(ka, kb); std::swap(va, vb);
    }
    voxel_heads_kernel<<<b, 1024, 0, st>>>(ka, n, w.starts, d_nvox, w.bad);
    PCP_CHECK_LAUNCH();
    voxel_reduce_kernel<<<dim3(gx, b), 256, 0, st>>>(d_x, n, va, w.starts, d_nvox, d_out_xyz, d_out_count);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

extern "C" int pcp_voxel_compact_f32(const float *d_x, int b, int64_t n, const float *h_origin, float voxel, int64_t capacity, float *d_out_xyz,
                                     int32_t *d_out_count, int32_t *d_nvox, void *d_workspace, size_t workspace_bytes, pcp_stream stream)
{
    PCP_REQUIRE(d_x && d_out_xyz && d_out_count && d_nvox && d_workspace && b >= 0 && n >= 1 && capacity >= 1, PCP_ERR_INVALID,
                "pcp_voxel_compact_f32: bad arguments");
    PCP_REQUIRE(voxel > 0.f, PCP_ERR_INVALID, "pcp_voxel_compact_f32: voxel size must be > 0");
    PCP_REQUIRE(n < 0xFFFFFFFFll && b <= 65535, PCP_ERR_UNSUPPORTED, "pcp_voxel_compact_f32: n or b too large");
    if (b == 0) return PCP_OK;
    VoxelWs w = carve(d_workspace, b, n);
    PCP_REQUIRE(w.bytes <= workspace_bytes, PCP_ERR_INVALID, "pcp_voxel_compact_f32: workspace too small (%zu < %zu)", workspace_bytes, w.bytes);
    cudaStream_t st = as_stream(stream);
    const float ox = h_origin ? h_origin[0] : 0.f, oy = h_origin ? h_origin[1] : 0.f, oz = h_origin ? h_origin[2] : 0.f;
    PCP_CUDA_OK(cudaMemsetAsync(w.bad, 0, (size_t)b * 4, st));
    PCP_CUDA_OK(cudaMemsetAsync(w.origin_enc, 0xFF, (size_t)b * 12, st));
    PCP_CUDA_OK(cudaMemsetAsync(w.max_enc, 0, (size_t)b * 12, st));
    PCP_CUDA_OK(cudaMemsetAsync(w.need_sort, 0, 4, st));
    const int vec_ok = ((reinterpret_cast<uintptr_t>(d_x) & 15u) == 0) && (n % 4 == 0);
    static int occ_minmax = 0, occ_clear = 0, occ_mark = 0, occ_accum = 0, occ_fin = 0, occ_zero = 0;
    const int64_t groups = vec_ok ? (n + 3) / 4 : n;
    voxel_minmax_kernel<<<dim3(wave_gx(voxel_minmax_kernel, b, vec_ok ? (n * 3 / 4 + 2) / 3 : n, &occ_minmax), b), 256, 0, st>>>(d_x, n, w.origin_enc, w.max_enc, vec_ok);
    PCP_CHECK_LAUNCH();
    voxel_grid_kernel<<<(b + 127) / 128, 128, 0, st>>>(w.origin_enc, w.max_enc, b, ox, oy, oz, h_origin ? 0 : 1, voxel, w.cap_cells, w.grid, w.need_sort);
    PCP_CHECK_LAUNCH();
    voxel_clear_kernel<<<dim3(wave_gx(voxel_clear_kernel, b, (w.cap_cells >> 7) + 1, &occ_clear), b), 256, 0, st>>>(w.grid, w.cap_cells, w.bitmap, w.need_sort);
    PCP_CHECK_LAUNCH();
    voxel_mark_kernel<<<dim3(wave_gx(voxel_mark_kernel, b, groups, &occ_mark), b), 256, 0, st>>>(d_x, n, w.grid, voxel, w.cap_cells, w.bitmap, w.bad, w.need_sort, vec_ok);
    PCP_CHECK_LAUNCH();
    voxel_rank_kernel<<<b, 1024, 0, st>>>(w.bitmap, w.grid, w.cap_cells, n, w.word_rank, d_nvox, w.bad, w.sums, w.cnt, w.need_sort);
    PCP_CHECK_LAUNCH();
    voxel_zero_kernel<<<dim3(wave_gx(voxel_zero_kernel, b, n, &occ_zero), b), 256, 0, st>>>(n, d_nvox, w.sums, w.need_sort);
    PCP_CHECK_LAUNCH();
    voxel_accum_kernel<<<dim3(wave_gx(voxel_accum_kernel, b, groups, &occ_accum), b), 256, 0, st>>>(d_x, n, w.grid, voxel, w.cap_cells, w.bitmap, w.word_rank, w.sums, w.cnt, w.need_sort, vec_ok);
    PCP_CHECK_LAUNCH();
    voxel_finalize_compact_kernel<<<dim3(wave_gx(voxel_finalize_compact_kernel, b, std::min<int64_t>(n, capacity), &occ_fin), b), 256, 0, st>>>(
        n, capacity, d_nvox, w.sums, d_out_xyz, d_out_count, w.need_sort);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}
