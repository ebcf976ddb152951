// voxel.cu — voxel-grid downsampling (north_star operator; no reference implementation, the
// contract is oracle/sampling.py:voxel_downsample — SURVEY.md §8 a10).
//
// Pipeline (all clouds in one launch each, everything asynchronous on the caller's stream):
//   1. per-cloud minimum corner (when no origin is given)             12 B/point read
//   2. key = floor((p - origin) / voxel) per axis -> 63-bit key, payload = point index
//   3. stable LSD radix sort of (key, index) per cloud, 8 passes x 8 bits, warp match-any ranking
//   4. run heads -> voxel ids (per-cloud scan) -> one thread per voxel sums its run in double
// Double accumulation makes the centroid independent of summation order (bit-stable results).
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

struct VoxelWs {
    unsigned long long *keys_a, *keys_b;
    unsigned *vals_a, *vals_b, *hist, *starts, *origin_enc;
    int *bad;
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
