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
voxel_minmax_kernel(const float *__restrict__ x, int64_t n, unsigned *__restrict__ min_enc, unsigned *__restrict__ max_enc)
{
    const int64_t cloud = blockIdx.y;
    const float *xc = x + cloud * n * 3;
    const float inf = __int_as_float(0x7f800000);
    float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { const float v = xc[3 * i + a]; lo[a] = fminf(lo[a], v); hi[a] = fmaxf(hi[a], v); }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(min_enc + cloud * 3 + a, enc_f32(lo[a]));
            atomicMax(max_enc + cloud * 3 + a, enc_f32(hi[a]));
        }
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

__global__ void __launch_bounds__(256)
voxel_mark_kernel(const float *__restrict__ x, int64_t n, const VoxelGrid *__restrict__ grid, float voxel, long long cap_cells,
                  unsigned *__restrict__ bitmap, int *__restrict__ bad)
{
    const int64_t cloud = blockIdx.y;
    const VoxelGrid g = grid[cloud];
    const float *xc = x + cloud * n * 3;
    unsigned *bm = bitmap + cloud * (cap_cells >> 5);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float p[3] = {xc[3 * i], xc[3 * i + 1], xc[3 * i + 2]};
        bool finite;
        const long long cell = voxel_cell(p, g, voxel, finite);
        if (!finite) { bad[cloud] = 1; continue; }
        atomicOr(bm + (cell >> 5), 1u << (cell & 31));
    }
}

// one CTA per cloud: exclusive popcount scan over the bitmap words -> rank of the first cell of every word; voxel count;
// accumulators of the occupied voxels zeroed
__global__ void __launch_bounds__(1024)
voxel_rank_kernel(const unsigned *__restrict__ bitmap, const VoxelGrid *__restrict__ grid, long long cap_cells, int64_t n,
                  unsigned *__restrict__ word_rank, int32_t *__restrict__ nvox, const int *__restrict__ bad, double *__restrict__ sums,
                  int *__restrict__ cnt)
{
    __shared__ unsigned part[1024];
    const int64_t cloud = blockIdx.x;
    const long long words = (grid[cloud].cells + 31) >> 5;
    const unsigned *bm = bitmap + cloud * (cap_cells >> 5);
    unsigned *wr = word_rank + cloud * (cap_cells >> 5);
    const long long per = (words + 1023) / 1024;
    const long long beg = threadIdx.x * per, end = min(words, beg + per);
    unsigned s = 0;
    for (long long w = beg; w < end; ++w) s += __popc(bm[w]);
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        unsigned v = threadIdx.x >= o ? part[threadIdx.x - o] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned r = part[threadIdx.x] - s;
    for (long long w = beg; w < end; ++w) { wr[w] = r; r += __popc(bm[w]); }
    const unsigned total = part[1023];
    if (threadIdx.x == 0) nvox[cloud] = bad[cloud] ? -1 : (int32_t)total;
    for (unsigned v = threadIdx.x; v < total; v += 1024) {
        double *sp = sums + (cloud * n + v) * 3;
        sp[0] = 0.0; sp[1] = 0.0; sp[2] = 0.0;
        cnt[cloud * n + v] = 0;
    }
}

__global__ void __launch_bounds__(256)
voxel_accum_kernel(const float *__restrict__ x, int64_t n, const VoxelGrid *__restrict__ grid, float voxel, long long cap_cells,
                   const unsigned *__restrict__ bitmap, const unsigned *__restrict__ word_rank, double *__restrict__ sums, int *__restrict__ cnt)
{
    const int64_t cloud = blockIdx.y;
    const VoxelGrid g = grid[cloud];
    const float *xc = x + cloud * n * 3;
    const unsigned *bm = bitmap + cloud * (cap_cells >> 5), *wr = word_rank + cloud * (cap_cells >> 5);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float p[3] = {xc[3 * i], xc[3 * i + 1], xc[3 * i + 2]};
        bool finite;
        const long long cell = voxel_cell(p, g, voxel, finite);
        if (!finite) continue;
        const unsigned w = __ldg(bm + (cell >> 5));
        const int64_t v = cloud * n + __ldg(wr + (cell >> 5)) + __popc(w & ((1u << (cell & 31)) - 1u));
        atomicAdd(sums + v * 3, (double)p[0]); atomicAdd(sums + v * 3 + 1, (double)p[1]); atomicAdd(sums + v * 3 + 2, (double)p[2]);
        atomicAdd(cnt + v, 1);
    }
}

__global__ void __launch_bounds__(256)
voxel_finalize_kernel(int64_t n, const int32_t *__restrict__ nvox, const double *__restrict__ sums, const int *__restrict__ cnt,
                      float *__restrict__ out_xyz, int32_t *__restrict__ out_count)
{
    const int64_t cloud = blockIdx.y;
    const int nv = max(nvox[cloud], 0);
    for (int64_t v = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; v < n; v += (int64_t)gridDim.x * blockDim.x) {
        float cx = 0.f, cy = 0.f, cz = 0.f;
        int c = 0;
        if (v < nv) {
            const double *sp = sums + (cloud * n + v) * 3;
            c = cnt[cloud * n + v];
            cx = (float)(sp[0] / (double)c); cy = (float)(sp[1] / (double)c); cz = (float)(sp[2] / (double)c);
        }
        float *o = out_xyz + (cloud * n + v) * 3;
        o[0] = cx; o[1] = cy; o[2] = cz;
        out_count[cloud * n + v] = c;
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
    w.sums = reinterpret_cast<double *>(take(bn * 24));
    w.cnt = reinterpret_cast<int *>(take(bn * 4));
    w.need_sort = reinterpret_cast<int *>(take(4));
    w.bytes = (size_t)(p - reinterpret_cast<char *>(base));
    return w;
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
        voxel_minmax_kernel<<<dim3(gx, b), 256, 0, st>>>(d_x, n, w.origin_enc, w.max_enc);
        PCP_CHECK_LAUNCH();
        voxel_grid_kernel<<<(b + 127) / 128, 128, 0, st>>>(w.origin_enc, w.max_enc, b, ox, oy, oz, h_origin ? 0 : 1, voxel, w.cap_cells, w.grid,
                                                          w.need_sort);
        PCP_CHECK_LAUNCH();
        int need_sort = 0;
        PCP_CUDA_OK(cudaMemcpyAsync(&need_sort, w.need_sort, 4, cudaMemcpyDeviceToHost, st));
        PCP_CUDA_OK(cudaStreamSynchronize(st));
        if (!need_sort) {
            PCP_CUDA_OK(cudaMemsetAsync(w.bitmap, 0, (size_t)b * (size_t)(w.cap_cells >> 3), st));
            voxel_mark_kernel<<<dim3(gx, b), 256, 0, st>>>(d_x, n, w.grid, voxel, w.cap_cells, w.bitmap, w.bad);
            PCP_CHECK_LAUNCH();
            voxel_rank_kernel<<<b, 1024, 0, st>>>(w.bitmap, w.grid, w.cap_cells, n, w.word_rank, d_nvox, w.bad, w.sums, w.cnt);
            PCP_CHECK_LAUNCH();
            voxel_accum_kernel<<<dim3(gx, b), 256, 0, st>>>(d_x, n, w.grid, voxel, w.cap_cells, w.bitmap, w.word_rank, w.sums, w.cnt);
            PCP_CHECK_LAUNCH();
            voxel_finalize_kernel<<<dim3(gx, b), 256, 0, st>>>(n, d_nvox, w.sums, w.cnt, d_out_xyz, d_out_count);
            PCP_CHECK_LAUNCH();
            return PCP_OK;
        }
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
