// tc_train.cu — tensor-core pieces of the training step (PCP_PRECISION_BF16): bf16 operands, fp32 accumulation in TMEM.
//   * pack_weight_image: fp32 weights in the flat parameter buffer -> the packed SWIZZLE_128B bf16 image tc_gemm streams
//     (forward: W as (K=cin x N=cout); input gradient: W^T as (K=cout x N=cin)).  Runs every step: weights change.
//   * tc_wgrad_kernel: the weight gradient dW (cin x cout) += A^T (cin x rows) . dZ (rows x cout).  Both operands are
//     row-major activations, i.e. MN-MAJOR for the MMA (the contraction index `rows` is the slow one), so the kernel
//     uses tcgen05's MN-major SWIZZLE_128B operand form: producer warps convert fp32 -> bf16 and write 16-byte chunks
//     (8 consecutive channels of one row) straight into the canonical ((8,n),(8,k)):((1,LBO),(8,SBO)) layout — no
//     transpose anywhere.  The reduction over rows is split across CTAs (split-K) and finished with fp32 atomics into the
//     zero-initialised flat gradient buffer.
#include <cstring>
#include <vector>

#include "tc_gemm.cuh"
#include "tc_train.cuh"
#include "umma.cuh"

namespace pcp {

using namespace umma;

// ------------------------------------------------------------------------------------------------
// weight image packer: element (k, n) of the logical (K x N) matrix is W[k*ldw + n] (transpose = 0) or W[n*ldw + k]
// (transpose = 1); tiles [nt][kt] of (bnt rows x 64 K) bf16, K-major SWIZZLE_128B (tc_gemm.cu:pack_gemm_b_image)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pack_weight_image_body(const float *__restrict__ W, int K, int N, int ldw, int transpose, int bnt,
                                                       const float *__restrict__ colscale, const float *__restrict__ sgn,
                                                       uint16_t *__restrict__ img, int64_t first, int64_t stride)
{
    const int KT = K >> 6, ntiles = (N + bnt - 1) / bnt;
    const int64_t total = (int64_t)ntiles * KT * bnt * 64;
    for (int64_t i = first; i < total; i += stride) {
        // decode i as (tile, r, k) with k fastest so that global reads of the non-transposed case are strided by ldw but
        // the writes stay inside one 128-byte swizzle row
        const int k = (int)(i & 63);
        const int64_t t = i >> 6;
        const int r = (int)(t % bnt);
        const int64_t tile = t / bnt;
        const int kt = (int)(tile % KT), nt = (int)(tile / KT);
        const int n = nt * bnt + r, kk = kt * 64 + k;
        float v = 0.f;
        if (n < N) v = transpose ? __ldg(W + (int64_t)n * ldw + kk) : __ldg(W + (int64_t)kk * ldw + n);
        if (colscale && n < N) v *= __ldg(colscale + n);
        if (sgn && n < N && !(__ldg(sgn + n) >= 0.f)) v = -v;
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        img[tile * (bnt * 64) + sw128_offset((uint32_t)bnt, (uint32_t)r, (uint32_t)k) / 2] = *reinterpret_cast<const uint16_t *>(&h);
    }
}

__global__ void __launch_bounds__(256)
pack_weight_image_kernel(const float *__restrict__ W, int K, int N, int ldw, int transpose, int bnt, const float *__restrict__ colscale,
                         const float *__restrict__ sgn, uint16_t *__restrict__ img)
{
    pack_weight_image_body(W, K, N, ldw, transpose, bnt, colscale, sgn, img, blockIdx.x * (int64_t)blockDim.x + threadIdx.x, (int64_t)gridDim.x * blockDim.x);
}

// every job of a table in one launch (blockIdx.y = job): the training step re-packs all its weight images once per step
__global__ void __launch_bounds__(256)
pack_weight_images_kernel(const PackJob *__restrict__ jobs)
{
    const PackJob j = jobs[blockIdx.y];
    pack_weight_image_body(j.W, j.K, j.N, j.ldw, j.transpose, j.bnt, nullptr, j.sgn, j.img, blockIdx.x * (int64_t)blockDim.x + threadIdx.x,
                           (int64_t)gridDim.x * blockDim.x);
}

int pack_weight_images(const PackJob *d_jobs, int njobs, cudaStream_t st)
{
    if (njobs <= 0) return PCP_OK;
    PCP_REQUIRE(njobs <= 65535, PCP_ERR_UNSUPPORTED, "pack_weight_images: too many jobs");
    // a (256 x 4096) image is 1 M elements: 256 x 256 threads take four each, blocks past a small image's end leave at once
    pack_weight_images_kernel<<<dim3(256, njobs), 256, 0, st>>>(d_jobs);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

int pack_weight_image(const float *d_W, int K, int N, int ldw, bool transpose, uint8_t *d_img, cudaStream_t st, const float *d_colscale,
                      const float *d_sgn)
{
    PCP_REQUIRE(K >= 64 && K % 64 == 0, PCP_ERR_UNSUPPORTED, "pack_weight_image: K=%d must be a multiple of 64", K);
    const int bnt = tc_gemm_bn_tile(N);
    const int64_t total = (int64_t)((N + bnt - 1) / bnt) * (K / 64) * bnt * 64;
    const int grid = (int)std::min<int64_t>((total + 255) / 256, (int64_t)sm_count() * 8);
    pack_weight_image_kernel<<<grid, 256, 0, st>>>(d_W, K, N, ldw, transpose ? 1 : 0, bnt, d_colscale, d_sgn, reinterpret_cast<uint16_t *>(d_img));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// ------------------------------------------------------------------------------------------------
// wgrad
// ------------------------------------------------------------------------------------------------
constexpr int kWgThreads = 288;                 // warps 0-7 producers (+ epilogue on warps 0-3), warp 8 MMA issuer
constexpr int kWgStages = 4;
constexpr uint32_t kWgTile = 16384;             // one operand tile: 128 (MN) x 64 (K = rows) bf16
constexpr uint32_t kWgStageBytes = 2 * kWgTile;
constexpr uint32_t kWgSmem = kWgStages * kWgStageBytes + 1024;
constexpr uint32_t kWgSBO = 1024;               // stride between groups of 8 rows (K)
constexpr uint32_t kWgLBO = 8192;               // stride between groups of 64 channels (MN)

// MN-major SWIZZLE_128B shared-memory descriptor
__device__ __forceinline__ uint64_t make_sdesc_mn(uint32_t saddr)
{
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
    d |= (uint64_t)(kWgLBO >> 4) << 16;
    d |= (uint64_t)(kWgSBO >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

struct WgradParams {
    const float *A; int64_t lda;          // (rows x cin) fp32 row-major
    const float *Z; int64_t ldz;          // (rows x cout) fp32 row-major (dZ)
    float *dW; int64_t lddw;              // (cin x cout) fp32, accumulated with atomics
    int64_t rows;
    int cin, cout;
    int64_t rows_per_split;               // multiple of 64
    int64_t split_stride;                 // 0: every split accumulates into dW with atomics; else split z STORES its partial at dW + z * stride
};

__global__ void __launch_bounds__(kWgThreads, 1)
tc_wgrad_kernel(const WgradParams p)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full[kWgStages], empty[kWgStages], acc_full;
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n0 = blockIdx.x * 128, m0 = blockIdx.y * 128;
    const int64_t r_begin = (int64_t)blockIdx.z * p.rows_per_split;
    const int64_t r_end = min(p.rows, r_begin + p.rows_per_split);
    const int KT = (int)((r_end - r_begin + 63) >> 6);

    if (tid == 0) {
        for (int i = 0; i < kWgStages; ++i) { mbar_init(&full[i], 8); mbar_init(&empty[i], 1); }
        mbar_init(&acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 8) tmem_alloc(&tmem_base, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    if (KT <= 0) {                                   // empty split: nothing to add (uniform per CTA)
        __syncthreads();
        if (warp == 8) tmem_dealloc(tmem, 128);
        return;
    }

    if (warp < 8) {
        // ---------------- producers: 4 A chunks + 4 Z chunks (16 B = 8 channels of one row) per thread per k-tile ----------------
        // chunk c = tid + 256*j : row k = c / 16 (0..63), channel group q = c % 16 (channels q*8 .. q*8+7)
        int krow[4], q[4];
        uint32_t soff[4];
        bool a_ok[4], z_ok[4];
        const float *ap[4], *zp[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = tid + 256 * j;
            krow[j] = c >> 4; q[j] = c & 15;
            soff[j] = (uint32_t)(q[j] >> 3) * kWgLBO + (uint32_t)(krow[j] >> 3) * kWgSBO + (uint32_t)(krow[j] & 7) * 128u +
                      (uint32_t)(((q[j] & 7) ^ (krow[j] & 7)) << 4);
            const int ma = m0 + q[j] * 8, nz = n0 + q[j] * 8;
            a_ok[j] = ma + 8 <= p.cin; z_ok[j] = nz + 8 <= p.cout;
            ap[j] = p.A + (a_ok[j] ? ma : 0);            // clamped column: loads stay unconditional, invalid chunks are zeroed
            zp[j] = p.Z + (z_ok[j] ? nz : 0);
        }
        // two k-tiles of loads in flight per thread (2 x 16 float4): the loads of tile kt + 2 are issued before tile kt is
        // converted, so a tile's loads have two full iterations to land
        float4 ra0[8], rz0[8], ra1[8], rz1[8];
        auto issue = [&](int kt, float4 (&ra)[8], float4 (&rz)[8]) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t r = min(r_begin + (int64_t)kt * 64 + krow[j], p.rows - 1);     // clamped row, zeroed below
                const float4 *sa = reinterpret_cast<const float4 *>(ap[j] + r * p.lda);
                const float4 *sz = reinterpret_cast<const float4 *>(zp[j] + r * p.ldz);
                ra[2 * j] = __ldg(sa); ra[2 * j + 1] = __ldg(sa + 1);
                rz[2 * j] = __ldg(sz); rz[2 * j + 1] = __ldg(sz + 1);
            }
        };
        auto publish = [&](int kt, const float4 (&ra)[8], const float4 (&rz)[8]) {
            const int s = kt % kWgStages;
            uint4 va[4], vz[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const bool row_ok = r_begin + (int64_t)kt * 64 + krow[j] < r_end;
                const float4 a0 = ra[2 * j], a1 = ra[2 * j + 1], z0 = rz[2 * j], z1 = rz[2 * j + 1];
                va[j] = (row_ok && a_ok[j]) ? make_uint4(pack_bf16x2(a0.x, a0.y), pack_bf16x2(a0.z, a0.w), pack_bf16x2(a1.x, a1.y), pack_bf16x2(a1.z, a1.w))
                                            : make_uint4(0u, 0u, 0u, 0u);
                vz[j] = (row_ok && z_ok[j]) ? make_uint4(pack_bf16x2(z0.x, z0.y), pack_bf16x2(z0.z, z0.w), pack_bf16x2(z1.x, z1.y), pack_bf16x2(z1.z, z1.w))
                                            : make_uint4(0u, 0u, 0u, 0u);
            }
            if (kt >= kWgStages) mbar_wait(&empty[s], (uint32_t)(((kt / kWgStages) - 1) & 1), 50);
            uint8_t *As = base + s * kWgStageBytes, *Zs = As + kWgTile;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                *reinterpret_cast<uint4 *>(As + soff[j]) = va[j];
                *reinterpret_cast<uint4 *>(Zs + soff[j]) = vz[j];
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
        };
        issue(0, ra0, rz0);
        if (KT > 1) issue(1, ra1, rz1);
        for (int kt = 0; kt < KT; kt += 2) {
            publish(kt, ra0, rz0);
            if (kt + 2 < KT) issue(kt + 2, ra0, rz0);
            if (kt + 1 < KT) {
                publish(kt + 1, ra1, rz1);
                if (kt + 3 < KT) issue(kt + 3, ra1, rz1);
            }
        }
        // ---------------- epilogue (warps 0-3): TMEM lane = input channel m, columns = output channels ----------------
        // Each lane holds 32 consecutive columns of ITS row; the 32 x 32 block goes through a warp-private shared tile so
        // that one atomic instruction covers four rows x 128 contiguous bytes (float4 red) instead of 32 rows x 4 bytes.
        if (warp < 4) {
            mbar_wait(&acc_full, 0, 51);
            tc_fence_after();
            float *stage = reinterpret_cast<float *>(base) + warp * (32 * 36);          // operand stages are idle now
            const bool vec_ok = (p.lddw & 3) == 0 && (((uintptr_t)p.dW & 15u) == 0);
            for (int c0 = 0; c0 < 128; c0 += 32) {
                if (n0 + c0 >= p.cout) break;
                uint32_t v[32];
                tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 32; e += 4)
                    *reinterpret_cast<uint4 *>(stage + lane * 36 + e) = make_uint4(v[e], v[e + 1], v[e + 2], v[e + 3]);
                __syncwarp();
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = 4 * i + (lane >> 3), c = (lane & 7) * 4;
                    const int m = m0 + warp * 32 + r, n = n0 + c0 + c;
                    const float4 q = *reinterpret_cast<const float4 *>(stage + r * 36 + c);
                    if (m < p.cin && n < p.cout) {               // cout % 8 == 0: a float4 never straddles the edge
                        float *dst = p.dW + (int64_t)blockIdx.z * p.split_stride + (int64_t)m * p.lddw + n;
                        if (p.split_stride) {                    // deterministic mode: this CTA is the only writer of its slice
                            dst[0] = q.x; dst[1] = q.y; dst[2] = q.z; dst[3] = q.w;
                        } else if (vec_ok) {
                            atomicAdd(reinterpret_cast<float4 *>(dst), q);
                        } else {
                            atomicAdd(dst, q.x); atomicAdd(dst + 1, q.y); atomicAdd(dst + 2, q.z); atomicAdd(dst + 3, q.w);
                        }
                    }
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------- MMA issuer: D[128 cin, 128 cout] += A^T[128, 16] . Z[16, 128] per instruction ----------------
        const uint32_t idesc = make_idesc_bf16(128, 128) | (1u << 15) | (1u << 16);      // a_major = b_major = MN
        for (int kt = 0; kt < KT; ++kt) {
            const int s = kt % kWgStages;
            mbar_wait(&full[s], (uint32_t)((kt / kWgStages) & 1), 52);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(base) + s * kWgStageBytes, z_addr = a_addr + kWgTile;
            if (elect_one_sync()) {
#pragma unroll
                for (int k = 0; k < 4; ++k)        // 16 rows (two 8-row groups) per MMA
                    mma_bf16_ss(tmem, make_sdesc_mn(a_addr + k * 2 * kWgSBO), make_sdesc_mn(z_addr + k * 2 * kWgSBO), idesc, (kt | k) > 0);
                mma_commit(&empty[s]);
                if (kt == KT - 1) mma_commit(&acc_full);
            }
            __syncwarp();
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem, 128);
}

// parts != nullptr: deterministic form — split z stores its partial product at parts + z * cin * lddw (nothing is accumulated with
// atomics) and *nparts receives the number of splits; the caller sums the slices in a fixed order
static int tc_wgrad_impl(const float *A, int64_t lda, int64_t rows, int cin, const float *dZ, int64_t ldz, int cout, float *dW, int64_t lddw,
                         float *parts, size_t parts_floats, int *nparts, cudaStream_t st)
{
    PCP_REQUIRE(A && dZ && (dW || parts), PCP_ERR_INVALID, "tc_wgrad: null operand");
    PCP_REQUIRE(cin % 8 == 0 && cout % 8 == 0 && lda % 4 == 0 && ldz % 4 == 0, PCP_ERR_UNSUPPORTED,
                "tc_wgrad: cin=%d / cout=%d must be multiples of 8 and the leading dimensions multiples of 4", cin, cout);
    PCP_REQUIRE((((uintptr_t)A | (uintptr_t)dZ) & 15u) == 0, PCP_ERR_UNSUPPORTED, "tc_wgrad: operands must be 16-byte aligned");
    if (rows <= 0) { if (nparts) *nparts = 0; return PCP_OK; }
    WgradParams p;
    p.A = A; p.lda = lda; p.Z = dZ; p.ldz = ldz; p.dW = dW; p.lddw = lddw; p.rows = rows; p.cin = cin; p.cout = cout; p.split_stride = 0;
    const int mt = (cin + 127) / 128, nt = (cout + 127) / 128;
    const int64_t ktiles = (rows + 63) / 64;
    int64_t splits = std::max<int64_t>(1, (int64_t)sm_count() / ((int64_t)mt * nt));            // one CTA per SM
    splits = std::min<int64_t>(splits, std::max<int64_t>(1, ktiles / 4));                        // at least 4 k-tiles per split
    splits = std::min<int64_t>(splits, 65535);
    const int64_t kps = (ktiles + splits - 1) / splits;
    p.rows_per_split = kps * 64;
    splits = (ktiles + kps - 1) / kps;
    if (parts) {
        p.split_stride = (int64_t)cin * lddw;
        PCP_REQUIRE((size_t)splits * (size_t)p.split_stride <= parts_floats, PCP_ERR_INVALID, "tc_wgrad: partial-product scratch too small");
        p.dW = parts;
        *nparts = (int)splits;
    }
    static bool attr[64] = {};
    int dev = 0;
    PCP_CUDA_OK(cudaGetDevice(&dev));
    PCP_REQUIRE(dev >= 0 && dev < 64, PCP_ERR_UNSUPPORTED, "tc_wgrad: device ordinal %d", dev);
    if (!attr[dev]) {
        PCP_CUDA_OK(cudaFuncSetAttribute(tc_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWgSmem));
        attr[dev] = true;
    }
    tc_wgrad_kernel<<<dim3(nt, mt, (unsigned)splits), kWgThreads, kWgSmem, st>>>(p);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

int tc_wgrad(const float *A, int64_t lda, int64_t rows, int cin, const float *dZ, int64_t ldz, int cout, float *dW, int64_t lddw,
             cudaStream_t st)
{
    return tc_wgrad_impl(A, lda, rows, cin, dZ, ldz, cout, dW, lddw, nullptr, 0, nullptr, st);
}

__global__ void __launch_bounds__(256)
sum_parts_kernel(const float *__restrict__ parts, int nparts, int64_t stride, int64_t count, float *__restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    double acc = 0.0;
    for (int z = 0; z < nparts; ++z) acc += (double)parts[(int64_t)z * stride + i];      // fixed order: the result does not depend on scheduling
    out[i] = (float)acc;
}

size_t tc_gram_scratch_floats(int cin) { return (size_t)sm_count() * cin * cin; }

// G (cin x cin) = A^T A of the bf16-rounded A, DETERMINISTIC: run to run the same bits (the atomics of tc_wgrad's split-K make the
// plain form differ at 1e-6, which a bf16-rounded network amplifies into a 1 % loss difference between two runs of the same step)
int tc_gram_det(const float *A, int64_t lda, int64_t rows, int cin, float *G, float *scratch, size_t scratch_floats, cudaStream_t st)
{
    int nparts = 0;
    PCP_TRY(tc_wgrad_impl(A, lda, rows, cin, A, lda, cin, nullptr, cin, scratch, scratch_floats, &nparts, st));
    const int64_t count = (int64_t)cin * cin;
    if (nparts == 0) { PCP_CUDA_OK(cudaMemsetAsync(G, 0, (size_t)count * 4, st)); return PCP_OK; }
    sum_parts_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(scratch, nparts, count, count, G);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// ------------------------------------------------------------------------------------------------
// per-cloud column extrema of Z = A . W without storing Z (training max-pool, train_pool.inl)
// ------------------------------------------------------------------------------------------------
// One CTA per 128-row tile of a cloud.  The tile of A (128 rows x 128 K) is converted to bf16 ONCE and stays in shared
// memory; the 128-channel tiles of W stream through two stages.  The MMA is issued TRANSPOSED — W tile as the A operand
// (M = 128 channels), the activation tile as the B operand (N = 128 rows) — so the accumulator has channels on the TMEM
// lanes and rows along the columns: the epilogue thread that owns a channel reduces over the rows serially in registers
// (no shuffles), two 64-bit atomicMax per thread and channel tile merge the row tiles of a cloud.  Two accumulator
// buffers (2 x 128 TMEM columns) overlap the MMA of tile nt + 1 with the epilogue of tile nt.
constexpr int kPxThreads = 192;                 // warps 0-3 A producer + epilogue, warp 4 MMA issuer, warp 5 W loader
constexpr uint32_t kPxTile = 32768;             // 128 x 128 bf16: two SW128 k-tiles of 16 KB
constexpr uint32_t kPxSmem = 3 * kPxTile + 1024;

struct PoolExtParams {
    const float *A;                      // (b * n) x 128 fp32 row-major
    const uint8_t *Wimg;                 // pack_weight_image(W, 128, cout): [cout/128][2] tiles of 16 KB
    unsigned long long *ext_max;         // b x cout packed keys (tc_gemm.cuh), zero-initialised
    int n, cout, mtiles;
};

constexpr int kPeThreads = 320;                 // warps 0-7: A producer, then two epilogue warpgroups; warp 8 MMA issuer, warp 9 W loader

// The epilogue is a chain of TMEM round trips (tcgen05.ld of 32 columns, wait, 32 compare/selects, four times per 128-channel
// block) and was the kernel's critical path with one warpgroup: 8 blocks x 4 round trips per CTA behind one another, 165 us.
// Two warpgroups take the channel blocks alternately (block nt lives in accumulator nt & 1, warpgroup nt & 1 reduces it), so
// two blocks are being reduced while the MMA of the next one runs; the A tile is converted by all eight warps.
__global__ void __launch_bounds__(kPeThreads, 2)
tc_pool_extrema_kernel(const PoolExtParams p)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t a_full, w_full[2], w_empty[2], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *As = base, *Ws = base + kPxTile;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = blockIdx.x / p.mtiles, mt = blockIdx.x % p.mtiles;
    const int row0 = mt * 128;
    const int rows_valid = min(128, p.n - row0);
    const int NT = p.cout >> 7;

    if (tid == 0) {
        mbar_init(&a_full, 8);
        for (int i = 0; i < 2; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
        fence_barrier_init();
    }
    if (warp == 8) tmem_alloc(&tmem_base, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;

    if (warp < 8) {
        // ---------------- A tile: 8 threads per row (32 B fp32 -> 16 B bf16), 32 rows per pass, clamped rows ----------------
        const int chunk = tid & 7, rsub = tid >> 3;
        const float *Ag = p.A + ((int64_t)g * p.n + row0) * 128 + chunk * 8;
        float4 raw[16];
#pragma unroll
        for (int kt = 0; kt < 2; ++kt)
#pragma unroll
            for (int pass = 0; pass < 4; ++pass) {
                const int r = min(pass * 32 + rsub, rows_valid - 1);
                const float4 *src = reinterpret_cast<const float4 *>(Ag + (int64_t)r * 128 + kt * 64);
                raw[kt * 8 + 2 * pass] = __ldg(src); raw[kt * 8 + 2 * pass + 1] = __ldg(src + 1);
            }
#pragma unroll
        for (int kt = 0; kt < 2; ++kt)
#pragma unroll
            for (int pass = 0; pass < 4; ++pass) {
                const float4 f0 = raw[kt * 8 + 2 * pass], f1 = raw[kt * 8 + 2 * pass + 1];
                *reinterpret_cast<uint4 *>(As + kt * 16384 + sw128_offset(128, (uint32_t)(pass * 32 + rsub), (uint32_t)chunk * 8)) =
                    make_uint4(pack_bf16x2(f0.x, f0.y), pack_bf16x2(f0.z, f0.w), pack_bf16x2(f1.x, f1.y), pack_bf16x2(f1.z, f1.w));
            }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full);
        // ---------------- epilogue: thread = channel, serial reduction over the tile's rows ----------------
        const int wg = warp >> 2, ch = (warp & 3) * 32 + lane;         // warpgroup = accumulator buffer; TMEM lane = channel
        const uint32_t tacc = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(wg * 128);
        for (int nt = wg; nt < NT; nt += 2) {
            mbar_wait(&acc_full[wg], (uint32_t)((nt >> 1) & 1), 61);
            tc_fence_after();
            float mx = -INFINITY;
            int rmx = 0;
            for (int c0 = 0; c0 < rows_valid; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tacc + (uint32_t)c0, v);
                tmem_ld_wait();
                if (c0 + 32 <= rows_valid) {
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        const float f = __uint_as_float(v[e]);
                        if (f > mx) { mx = f; rmx = c0 + e; }
                    }
                } else {
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        const float f = __uint_as_float(v[e]);
                        if (c0 + e < rows_valid && f > mx) { mx = f; rmx = c0 + e; }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[wg]);
            const size_t o = (size_t)g * p.cout + (size_t)nt * 128 + ch;
            if (mx > -INFINITY)
                atomicMax(p.ext_max + o, ((unsigned long long)tc_float_key(__float_as_uint(mx)) << 32) | (0xffffffffu - (uint32_t)(row0 + rmx)));
        }
    } else if (warp == 8) {
        // ---------------- MMA issuer: D[channel, row] = W tile (128 ch x 128 K) . A tile^T ----------------
        const uint32_t idesc = make_idesc_bf16(128, 128);
        const uint32_t a_lo = (smem_u32(As) & 0x3FFFFu) >> 4, w_lo = (smem_u32(Ws) & 0x3FFFFu) >> 4;
        mbar_wait(&a_full, 0, 62);
        for (int nt = 0; nt < NT; ++nt) {
            const int s = nt & 1;
            mbar_wait(&w_full[s], (uint32_t)((nt >> 1) & 1), 63);
            if (nt >= 2) mbar_wait(&acc_empty[s], (uint32_t)(((nt >> 1) - 1) & 1), 64);
            tc_fence_after();
            if (elect_one_sync()) {
#pragma unroll
                for (int kt = 0; kt < 2; ++kt)
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        mma_bf16_ss(tmem + (uint32_t)(s * 128), sdesc_from_lo(w_lo + (uint32_t)s * (kPxTile >> 4) + kt * 1024 + k * 2),
                                    sdesc_from_lo(a_lo + kt * 1024 + k * 2), idesc, (kt | k) > 0);
                mma_commit(&w_empty[s]);
                mma_commit(&acc_full[s]);
            }
            __syncwarp();
        }
    } else {
        // ---------------- W loader ----------------
        if (lane == 0) {
            for (int nt = 0; nt < NT; ++nt) {
                const int s = nt & 1;
                if (nt >= 2) mbar_wait(&w_empty[s], (uint32_t)(((nt >> 1) - 1) & 1), 65);
                mbar_arrive_expect_tx(&w_full[s], kPxTile);
                bulk_g2s(Ws + s * kPxTile, p.Wimg + (size_t)nt * kPxTile, 16384, &w_full[s]);
                bulk_g2s(Ws + s * kPxTile + 16384, p.Wimg + (size_t)nt * kPxTile + 16384, 16384, &w_full[s]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 8) tmem_dealloc(tmem, 256);
}

int tc_pool_extrema(const float *A, int b, int64_t n, const uint8_t *Wimg, int cout, unsigned long long *ext_max, cudaStream_t st)
{
    PCP_REQUIRE(A && Wimg && ext_max, PCP_ERR_INVALID, "tc_pool_extrema: null operand");
    PCP_REQUIRE(cout >= 128 && cout % 128 == 0 && n >= 1 && n <= 0x7fffffff, PCP_ERR_UNSUPPORTED, "tc_pool_extrema: cout=%d must be a multiple of 128", cout);
    if (b <= 0) return PCP_OK;
    PoolExtParams p;
    p.A = A; p.Wimg = Wimg; p.ext_max = ext_max; p.n = (int)n; p.cout = cout; p.mtiles = (int)((n + 127) / 128);
    const int64_t grid = (int64_t)b * p.mtiles;
    PCP_REQUIRE(grid <= 0x7fffffff, PCP_ERR_UNSUPPORTED, "tc_pool_extrema: %lld row tiles exceed one launch", (long long)grid);
    static bool attr = false;
    if (!attr) {
        PCP_CUDA_OK(cudaFuncSetAttribute(tc_pool_extrema_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPxSmem));
        attr = true;
    }
    tc_pool_extrema_kernel<<<(unsigned)grid, kPeThreads, kPxSmem, st>>>(p);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}


// ------------------------------------------------------------------------------------------------
// fp32-contract inference on the bf16 tensor cores (PCP_PRECISION_FP32_TC): g[cloud, ch] = max over the cloud's rows of
// relu(A . W' + t) for the three max-pooled 128 -> 1024 layers (PointNet.py:425-429, :242-248), with the product evaluated as
// hi(W) hi(A) + lo(W) hi(A) + hi(W) lo(A)  (hi = bf16(x), lo = bf16(x - hi): ~16 mantissa bits, fp32 accumulation in TMEM).
// Same structure as tc_pool_extrema_kernel (transposed MMA: channels on the TMEM lanes, rows along the columns, the owning
// thread reduces over the rows in registers); BN is folded into W' / t, so relu(max z + t) is the pooled value whatever the
// sign of gamma; clouds' row tiles merge through atomicMax on the float bits (non-negative floats order like integers).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kP3Smem = 6 * kPxTile + 1024;     // A hi | A lo | 2 stages x (W hi | W lo)

struct PoolMax3Params {
    const float *A;                      // (b * n) x 128 fp32 row-major
    const uint8_t *Whi, *Wlo;            // pack_gemm_b_image(hi(W') / lo(W'), 128, cout): [cout/128][2] tiles of 16 KB
    const float *t;                      // cout: folded shift
    float *g;                            // b x cout, zero-initialised
    int n, cout, mtiles;
};

__global__ void __launch_bounds__(kPxThreads, 1)
tc_pool_max3_kernel(const PoolMax3Params p)
{
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t a_full, w_full[2], w_empty[2], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t *Ahi = base, *Alo = base + kPxTile, *Ws = base + 2 * kPxTile;      // stage s: Ws + s * 2 * kPxTile = hi | lo
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = blockIdx.x / p.mtiles, mt = blockIdx.x % p.mtiles;
    const int row0 = mt * 128;
    const int rows_valid = min(128, p.n - row0);
    const int NT = p.cout >> 7;

    if (tid == 0) {
        mbar_init(&a_full, 4);
        for (int i = 0; i < 2; ++i) { mbar_init(&w_full[i], 1); mbar_init(&w_empty[i], 1); mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&tmem_base, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;

    if (warp < 4) {
        const int chunk = tid & 7, rsub = tid >> 3;
        const float *Ag = p.A + ((int64_t)g * p.n + row0) * 128 + chunk * 8;
#pragma unroll
        for (int kt = 0; kt < 2; ++kt) {
            float4 raw[16];
#pragma unroll
            for (int pass = 0; pass < 8; ++pass) {
                const int r = min(pass * 16 + rsub, rows_valid - 1);
                const float4 *src = reinterpret_cast<const float4 *>(Ag + (int64_t)r * 128 + kt * 64);
                raw[2 * pass] = __ldg(src); raw[2 * pass + 1] = __ldg(src + 1);
            }
#pragma unroll
            for (int pass = 0; pass < 8; ++pass) {
                const float f[8] = {raw[2 * pass].x, raw[2 * pass].y, raw[2 * pass].z, raw[2 * pass].w,
                                    raw[2 * pass + 1].x, raw[2 * pass + 1].y, raw[2 * pass + 1].z, raw[2 * pass + 1].w};
                float l[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) l[e] = f[e] - __bfloat162float(__float2bfloat16_rn(f[e]));
                const uint32_t off = kt * 16384 + sw128_offset(128, (uint32_t)(pass * 16 + rsub), (uint32_t)chunk * 8);
                *reinterpret_cast<uint4 *>(Ahi + off) = make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
                *reinterpret_cast<uint4 *>(Alo + off) = make_uint4(pack_bf16x2(l[0], l[1]), pack_bf16x2(l[2], l[3]), pack_bf16x2(l[4], l[5]), pack_bf16x2(l[6], l[7]));
            }
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_full);
        for (int nt = 0; nt < NT; ++nt) {
            const int buf = nt & 1;
            mbar_wait(&acc_full[buf], (uint32_t)((nt >> 1) & 1), 66);
            tc_fence_after();
            float mx = -INFINITY;
            for (int c0 = 0; c0 < rows_valid; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(buf * 128 + c0), v);
                tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 32; ++e)
                    if (c0 + e < rows_valid) mx = fmaxf(mx, __uint_as_float(v[e]));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
            const int ch = nt * 128 + tid;
            const float y = fmaxf(mx + __ldg(p.t + ch), 0.f);
            if (y > 0.f) atomicMax(reinterpret_cast<unsigned *>(p.g + (size_t)g * p.cout + ch), __float_as_uint(y));
        }
    } else if (warp == 4) {
        const uint32_t idesc = make_idesc_bf16(128, 128);
        const uint32_t ahi = (smem_u32(Ahi) & 0x3FFFFu) >> 4, alo = (smem_u32(Alo) & 0x3FFFFu) >> 4, w0 = (smem_u32(Ws) & 0x3FFFFu) >> 4;
        mbar_wait(&a_full, 0, 67);
        for (int nt = 0; nt < NT; ++nt) {
            const int s = nt & 1;
            mbar_wait(&w_full[s], (uint32_t)((nt >> 1) & 1), 68);
            if (nt >= 2) mbar_wait(&acc_empty[s], (uint32_t)(((nt >> 1) - 1) & 1), 69);
            tc_fence_after();
            if (elect_one_sync()) {
                const uint32_t whi = w0 + (uint32_t)s * (2 * kPxTile >> 4), wlo = whi + (kPxTile >> 4);
                const uint32_t wsel[3] = {whi, wlo, whi}, asel[3] = {ahi, ahi, alo};
#pragma unroll
                for (int part = 0; part < 3; ++part)
#pragma unroll
                    for (int kt = 0; kt < 2; ++kt)
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            mma_bf16_ss(tmem + (uint32_t)(s * 128), sdesc_from_lo(wsel[part] + kt * 1024 + k * 2), sdesc_from_lo(asel[part] + kt * 1024 + k * 2),
                                        idesc, (part | kt | k) > 0);
                mma_commit(&w_empty[s]);
                mma_commit(&acc_full[s]);
            }
            __syncwarp();
        }
    } else {
        if (lane == 0) {
            for (int nt = 0; nt < NT; ++nt) {
                const int s = nt & 1;
                if (nt >= 2) mbar_wait(&w_empty[s], (uint32_t)(((nt >> 1) - 1) & 1), 70);
                mbar_arrive_expect_tx(&w_full[s], 2 * kPxTile);
                uint8_t *dst = Ws + (size_t)s * 2 * kPxTile;
                bulk_g2s(dst, p.Whi + (size_t)nt * kPxTile, 16384, &w_full[s]);
                bulk_g2s(dst + 16384, p.Whi + (size_t)nt * kPxTile + 16384, 16384, &w_full[s]);
                bulk_g2s(dst + kPxTile, p.Wlo + (size_t)nt * kPxTile, 16384, &w_full[s]);
                bulk_g2s(dst + kPxTile + 16384, p.Wlo + (size_t)nt * kPxTile + 16384, 16384, &w_full[s]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 256);
}

int tc_pool_max3(const float *A, int b, int64_t n, const uint8_t *Whi, const uint8_t *Wlo, const float *t, int cout, float *g, cudaStream_t st)
{
    PCP_REQUIRE(A && Whi && Wlo && t && g, PCP_ERR_INVALID, "tc_pool_max3: null operand");
    PCP_REQUIRE(cout >= 128 && cout % 128 == 0 && n >= 1 && n <= 0x7fffffff, PCP_ERR_UNSUPPORTED, "tc_pool_max3: cout=%d must be a multiple of 128", cout);
    if (b <= 0) return PCP_OK;
    PoolMax3Params p;
    p.A = A; p.Whi = Whi; p.Wlo = Wlo; p.t = t; p.g = g; p.n = (int)n; p.cout = cout; p.mtiles = (int)((n + 127) / 128);
    const int64_t grid = (int64_t)b * p.mtiles;
    PCP_REQUIRE(grid <= 0x7fffffff, PCP_ERR_UNSUPPORTED, "tc_pool_max3: %lld row tiles exceed one launch", (long long)grid);
    static bool attr = false;
    if (!attr) {
        PCP_CUDA_OK(cudaFuncSetAttribute(tc_pool_max3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kP3Smem));
        attr = true;
    }
    tc_pool_max3_kernel<<<(unsigned)grid, kPxThreads, kP3Smem, st>>>(p);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

}  // namespace pcp

using namespace pcp;

// test hook: dW (cin x cout) = A^T . dZ with bf16-rounded operands, host in / host out, synchronous
extern "C" int pcp_selftest_wgrad(const float *h_a, const float *h_dz, int64_t rows, int cin, int cout, float *h_dw)
{
    PCP_REQUIRE(h_a && h_dz && h_dw && rows >= 1 && cin >= 8 && cout >= 8, PCP_ERR_INVALID, "pcp_selftest_wgrad: bad arguments");
    float *da = nullptr, *dz = nullptr, *dw = nullptr;
    PCP_CUDA_OK(cudaMalloc(&da, (size_t)rows * cin * 4)); PCP_CUDA_OK(cudaMalloc(&dz, (size_t)rows * cout * 4));
    PCP_CUDA_OK(cudaMalloc(&dw, (size_t)cin * cout * 4));
    PCP_CUDA_OK(cudaMemcpy(da, h_a, (size_t)rows * cin * 4, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(dz, h_dz, (size_t)rows * cout * 4, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemset(dw, 0, (size_t)cin * cout * 4));
    int rc = tc_wgrad(da, cin, rows, cin, dz, cout, cout, dw, cout, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc == PCP_OK && e != cudaSuccess) {
        unsigned info[4] = {0, 0, 0, 0};
        cudaMemcpyFromSymbol(info, g_mbar_timeout_info, sizeof(info));
        set_error("pcp_selftest_wgrad: %s (timeout info %08x %u %u %u)", cudaGetErrorString(e), info[0], info[1], info[2], info[3]);
        rc = PCP_ERR_CUDA;
    }
    if (rc == PCP_OK) PCP_CUDA_OK(cudaMemcpy(h_dw, dw, (size_t)cin * cout * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(dz); cudaFree(dw);
    return rc;
}

// test hook: C (rows x n) (+)= A (rows x k) . W with the device-packed weight image; W is (k x n) row-major, or (n x k)
// row-major when transpose_w (the input-gradient form dA = dZ . W^T); host in / host out, synchronous
static int selftest_tc_mm(const float *h_a, const float *h_w, int64_t rows, int k, int n, int transpose_w, int accumulate, int ksplit, float *h_c)
{
    using namespace pcp;
    PCP_REQUIRE(h_a && h_w && h_c && rows >= 1 && k >= 64 && n >= 1 && ksplit >= 1, PCP_ERR_INVALID, "pcp_selftest_tc_mm: bad arguments");
    float *da = nullptr, *dw = nullptr, *dc = nullptr;
    uint8_t *img = nullptr;
    PCP_CUDA_OK(cudaMalloc(&da, (size_t)rows * k * 4)); PCP_CUDA_OK(cudaMalloc(&dw, (size_t)k * n * 4));
    PCP_CUDA_OK(cudaMalloc(&dc, (size_t)rows * n * 4)); PCP_CUDA_OK(cudaMalloc(&img, gemm_b_image_bytes(k, n)));
    PCP_CUDA_OK(cudaMemcpy(da, h_a, (size_t)rows * k * 4, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(dw, h_w, (size_t)k * n * 4, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(dc, h_c, (size_t)rows * n * 4, cudaMemcpyHostToDevice));
    int rc = pack_weight_image(dw, k, n, transpose_w ? k : n, transpose_w != 0, img, nullptr);
    if (rc == PCP_OK) {
        TcGemmParams p;
        p.A = da; p.a_f32 = 1; p.lda = k; p.Bimg = img; p.C32 = dc; p.ldc = n; p.M = (int)rows; p.N = n; p.K = k;
        p.accumulate = accumulate ? 1 : 0;
        p.ksplit = ksplit;
        rc = tc_gemm(p, nullptr);
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (rc == PCP_OK && e != cudaSuccess) {
        set_error("pcp_selftest_tc_mm: %s", cudaGetErrorString(e));
        rc = PCP_ERR_CUDA;
    }
    if (rc == PCP_OK) PCP_CUDA_OK(cudaMemcpy(h_c, dc, (size_t)rows * n * 4, cudaMemcpyDeviceToHost));
    cudaFree(da); cudaFree(dw); cudaFree(dc); cudaFree(img);
    return rc;
}

extern "C" int pcp_selftest_tc_mm(const float *h_a, const float *h_w, int64_t rows, int k, int n, int transpose_w, int accumulate,
                                  float *h_c)
{
    return selftest_tc_mm(h_a, h_w, rows, k, n, transpose_w, accumulate, 1, h_c);
}

// the same with the k-tiles dealt over ksplit CTAs that ADD their products into h_c (TcGemmParams::ksplit; k / 64 a multiple of it)
extern "C" int pcp_selftest_tc_mm_splitk(const float *h_a, const float *h_w, int64_t rows, int k, int n, int transpose_w, int ksplit,
                                         float *h_c)
{
    return selftest_tc_mm(h_a, h_w, rows, k, n, transpose_w, 0, ksplit, h_c);
}

// development hook: time tc_gemm alone on device-resident synthetic operands (bf16 or fp32 activations, zero weights are fine:
// the kernel is memory-bound).  Buffers rotate through enough copies to exceed the 126 MB L2.  out_kind: 0 fp32, 1 bf16, 2 softmax.
extern "C" int pcp_bench_tc_gemm(int64_t rows, int k, int n, int out_kind, int a_f32, int relu, int iters, int dbg, float *h_ms)
{
    PCP_REQUIRE(h_ms && rows >= 1 && k >= 64 && n >= 1 && iters >= 1, PCP_ERR_INVALID, "pcp_bench_tc_gemm: bad arguments");
    const size_t a_bytes = (size_t)rows * k * (a_f32 ? 4 : 2), c_bytes = (size_t)rows * n * (out_kind == 1 ? 2 : 4);
    const int nbuf = (int)std::max<size_t>(2, (size_t)400e6 / (a_bytes + c_bytes) + 1);
    uint8_t *a = nullptr, *c = nullptr, *img = nullptr;
    float *bias = nullptr;
    PCP_CUDA_OK(cudaMalloc(&a, a_bytes * nbuf)); PCP_CUDA_OK(cudaMalloc(&c, c_bytes * nbuf));
    PCP_CUDA_OK(cudaMalloc(&img, gemm_b_image_bytes(k, n))); PCP_CUDA_OK(cudaMalloc(&bias, (size_t)n * 4));
    PCP_CUDA_OK(cudaMemset(a, 0, a_bytes * nbuf)); PCP_CUDA_OK(cudaMemset(img, 0, gemm_b_image_bytes(k, n))); PCP_CUDA_OK(cudaMemset(bias, 0, (size_t)n * 4));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int rc = PCP_OK;
    for (int it = -3; it < iters && rc == PCP_OK; ++it) {
        if (it == 0) cudaEventRecord(e0, nullptr);
        const int bsel = ((it % nbuf) + nbuf) % nbuf;
        TcGemmParams p;
        p.A = a + a_bytes * bsel; p.a_f32 = a_f32; p.lda = k; p.Bimg = img; p.bias = bias; p.ldc = n; p.M = (int)rows; p.N = n; p.K = k;
        p.relu = relu; p.dbg = dbg;
        if (out_kind == 0) p.C32 = reinterpret_cast<float *>(c + c_bytes * bsel);
        else if (out_kind == 1) p.C16 = reinterpret_cast<__nv_bfloat16 *>(c + c_bytes * bsel);
        else p.Csoftmax = reinterpret_cast<float *>(c + c_bytes * bsel);
        rc = tc_gemm(p, nullptr);
    }
    cudaEventRecord(e1, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc == PCP_OK && e != cudaSuccess) { set_error("pcp_bench_tc_gemm: %s", cudaGetErrorString(e)); rc = PCP_ERR_CUDA; }
    float ms = 0.f;
    if (rc == PCP_OK) { cudaEventElapsedTime(&ms, e0, e1); *h_ms = ms / iters; }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(a); cudaFree(c); cudaFree(img); cudaFree(bias);
    return rc;
}
