// tc_gemm.cu — generic bf16 tcgen05 GEMM: 128 x BNt output tile per CTA, K streamed in 64-element blocks
// through a 3-stage shared-memory ring.  A tiles are converted (fp32 -> bf16 if needed) and written by four
// producer warps straight into the K-major SWIZZLE_128B operand image; weight tiles arrive pre-packed by
// cp.async.bulk; one elected thread issues tcgen05.mma; the producer warps then run the epilogue
// (TMEM -> registers -> bias / ReLU / softmax -> global).  96 KB smem + 128 TMEM columns => 2 CTAs per SM.
#include <cstring>

#include "tc_gemm.cuh"
#include "umma.cuh"

namespace pcp {

using namespace umma;

constexpr int kGemmThreads = 192;     // warps 0-3 producer + epilogue, warp 4 MMA issuer, warp 5 weight loader
constexpr int kGemmStages = 3;
constexpr uint32_t kGemmStageBytes = 32768;   // 16 KB A + up to 16 KB B
constexpr uint32_t kGemmSmem = kGemmStages * kGemmStageBytes + 1024;

__device__ __forceinline__ float bf16_trunc_rn(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// kMinBlocks = 3: the K <= 128 per-point layers of the training step.  One k-tile or two means one or two ring stages
// (33 / 65 KB of shared memory), and at 96 registers three CTAs fit an SM: each CTA is one dependent chain (weights and
// activations in, one MMA group, TMEM out), so the tiles in flight per SM are what hides its latency.
// kSplitK: the k-tiles are dealt over grid.z CTAs which add their products with atomics (TcGemmParams::ksplit); a separate
// instantiation so that the plain kernel's code is exactly what it was (the fp32_tc forward lost 5 % to the shared version).
template <int kMinBlocks, bool kSplitK>
__global__ void __launch_bounds__(kGemmThreads, kMinBlocks)
tc_gemm_kernel(const TcGemmParams p, int bnt, int mtiles, int img_rows, int nsplit)
{   // bnt: MMA N of this CTA; img_rows: rows of a weight-image tile; nsplit = img_rows / bnt column blocks per image tile
    extern __shared__ uint8_t smem_raw[];
    __shared__ uint64_t full[kGemmStages], empty[kGemmStages], acc_full;
    __shared__ uint32_t tmem_base;
    uint8_t *base = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // A (img_rows x 64) SWIZZLE_128B tile is a stack of 1024-byte 8-row atoms, so rows [j bnt, (j + 1) bnt) of it are a
    // contiguous, valid bnt-row operand: skinny GEMMs (M = one row tile) split their 128-column tiles over several CTAs, each
    // streaming its share of the weights (4 CTAs pulling a 1 MB image had been the whole cost of a 1024 -> 512 head layer)
    const int ntile = blockIdx.x / nsplit, nsub = blockIdx.x % nsplit;
    const int ncol0 = ntile * img_rows + nsub * bnt;
    const int g = blockIdx.y / mtiles, mt = blockIdx.y % mtiles;
    const int row0 = mt * 128;
    const int rows_valid = min(128, p.M - row0);
    const int KT = kSplitK ? (p.K >> 6) / (int)gridDim.z : (p.K >> 6);     // this CTA's k-tiles: [kt0, kt0 + KT) of the matrix
    const int kt0 = kSplitK ? (int)blockIdx.z * KT : 0;
    const uint32_t b_tile_bytes = (uint32_t)bnt * 128u;

    if (tid == 0) {
        for (int i = 0; i < kGemmStages; ++i) { mbar_init(&full[i], 5); mbar_init(&empty[i], 1); }
        mbar_init(&acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&tmem_base, 128);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    // programmatic dependent launch: the skinny head GEMMs run back to back in grids of 1-32 CTAs, so the next launch's
    // scheduling + prologue (barrier init, TMEM allocation) is let loose now and overlaps this kernel's body; nothing of the
    // predecessor's output is touched before the wait (no-ops when the launch did not opt in)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (warp < 4) {
        // ---------------- A producer: 8 threads per row (16 B chunks), 16 rows per pass ----------------
        const int chunk = tid & 7, rsub = tid >> 3;
        const char *Abase = reinterpret_cast<const char *>(p.A);
        const size_t esz = p.a_f32 ? 4 : 2;
        const size_t a_off = ((size_t)g * p.a_group_stride + (size_t)row0 * p.lda + (size_t)kt0 * 64) * esz;
        // Rows past the end of the matrix are clamped to the last valid row instead of predicated: every load is
        // unconditional, so the compiler issues all loads of a k-tile back to back (ONE memory round trip per tile —
        // predicated loads had been serialised into eight dependent round trips, 3.7 us per k-tile); the duplicated
        // rows only produce output rows that the epilogue never stores.
        const char *rowp[8];
#pragma unroll
        for (int pass = 0; pass < 8; ++pass) {
            const int r = min(pass * 16 + rsub, rows_valid - 1);
            rowp[pass] = Abase + a_off + ((size_t)r * p.lda + (size_t)chunk * 8) * esz;
        }
        auto publish = [&](int kt, const uint4 (&v)[8]) {
            const int s = kt % kGemmStages;
            if (kt >= kGemmStages) mbar_wait(&empty[s], (uint32_t)(((kt / kGemmStages) - 1) & 1), 40);
            uint8_t *As = base + s * kGemmStageBytes;
#pragma unroll
            for (int pass = 0; pass < 8; ++pass)
                *reinterpret_cast<uint4 *>(As + sw128_offset(128, (uint32_t)(pass * 16 + rsub), (uint32_t)chunk * 8)) = v[pass];
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[s]);
        };
        if (p.a_f32) {
            // fp32 activations (pooled features): one k-tile (16 x float4) in flight, converted to bf16 on arrival
            float4 raw[16];
            const int KTs = p.a_split3 ? KT / 3 : KT;          // k-tiles of the source matrix (split3: every tile is read three times; never split-K)
            auto issue = [&](int kt) {
                const int ks = p.a_split3 ? kt % KTs : kt;
#pragma unroll
                for (int pass = 0; pass < 8; ++pass) {
                    const float4 *src = reinterpret_cast<const float4 *>(rowp[pass] + (size_t)ks * 256);
                    raw[2 * pass] = __ldg(src);
                    raw[2 * pass + 1] = __ldg(src + 1);
                }
            };
            issue(0);
            for (int kt = 0; kt < KT; ++kt) {
                uint4 v[8];
                const bool lo_part = p.a_split3 && kt >= 2 * KTs;
#pragma unroll
                for (int pass = 0; pass < 8; ++pass) {
                    float4 f0 = raw[2 * pass], f1 = raw[2 * pass + 1];
                    if (lo_part) {                             // lo(A) = bf16(A - float(bf16(A))): exact difference, then rounded
                        f0.x -= bf16_trunc_rn(f0.x); f0.y -= bf16_trunc_rn(f0.y); f0.z -= bf16_trunc_rn(f0.z); f0.w -= bf16_trunc_rn(f0.w);
                        f1.x -= bf16_trunc_rn(f1.x); f1.y -= bf16_trunc_rn(f1.y); f1.z -= bf16_trunc_rn(f1.z); f1.w -= bf16_trunc_rn(f1.w);
                    }
                    v[pass] = make_uint4(pack_bf16x2(f0.x, f0.y), pack_bf16x2(f0.z, f0.w), pack_bf16x2(f1.x, f1.y), pack_bf16x2(f1.z, f1.w));
                }
                if (kt + 1 < KT) issue(kt + 1);
                publish(kt, v);
            }
        } else {
            // bf16 activations: two k-tiles (2 x 8 x 16 B) in flight
            uint4 r0[8], r1[8];
            auto issue = [&](int kt, uint4 (&dst)[8]) {
#pragma unroll
                for (int pass = 0; pass < 8; ++pass) dst[pass] = __ldg(reinterpret_cast<const uint4 *>(rowp[pass] + (size_t)kt * 128));
            };
            const bool ld = !(p.dbg & 2);
            if (ld) issue(0, r0);
            if (ld && KT > 1) issue(1, r1);
            for (int kt = 0; kt < KT; kt += 2) {
                publish(kt, r0);
                if (ld && kt + 2 < KT) issue(kt + 2, r0);
                if (kt + 1 < KT) {
                    publish(kt + 1, r1);
                    if (ld && kt + 3 < KT) issue(kt + 3, r1);
                }
            }
        }
        // ---------------- epilogue: thread = output row ----------------
        mbar_wait(&acc_full, 0, 41);
        tc_fence_after();
        const int row = tid;
        const bool row_ok = row < rows_valid && !(p.dbg & 1);
        const float *bias = p.bias ? p.bias + (size_t)g * p.bias_group_stride : nullptr;
        const size_t c_off = (size_t)g * p.c_group_stride + (size_t)(row0 + row) * p.ldc;
        for (int c0 = 0; c0 < bnt; c0 += 32) {
            uint32_t v[32];
            tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
            tmem_ld_wait();
            float o[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                const int n = ncol0 + c0 + e;
                float f = __uint_as_float(v[e]) + ((bias && n < p.N) ? __ldg(bias + n) : 0.f);
                if (p.relu) f = f > 0.f ? f : 0.f;
                o[e] = f;
            }
            const int nbase = ncol0 + c0;
            if (p.C32) {
                float *Cg = p.C32 + (size_t)g * p.c_group_stride;
                if (p.dbg & 1) {
                } else if (kSplitK) {
                    if (row_ok) {
                        float *dst = p.C32 + c_off + nbase;
#pragma unroll
                        for (int e = 0; e < 32; ++e) if (nbase + e < p.N) atomicAdd(dst + e, o[e]);
                    }
                } else if (nbase + 32 <= p.N && (p.ldc & 3) == 0 && (((uintptr_t)Cg & 15u) == 0)) {
                    // warp-uniform fast path: the lane's 32 columns of ITS row go through a warp-private shared tile (the
                    // operand stages are idle once acc_full fired) so that one store instruction writes four rows x 128
                    // contiguous bytes instead of 16 bytes in each of 32 rows
                    float *stage = reinterpret_cast<float *>(base) + warp * (32 * 36);
#pragma unroll
                    for (int e = 0; e < 32; e += 4) *reinterpret_cast<float4 *>(stage + lane * 36 + e) = make_float4(o[e], o[e + 1], o[e + 2], o[e + 3]);
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int rr = warp * 32 + 4 * i + (lane >> 3), c = (lane & 7) * 4;
                        if (rr < rows_valid) {
                            float4 q = *reinterpret_cast<const float4 *>(stage + (4 * i + (lane >> 3)) * 36 + c);
                            float4 *dst = reinterpret_cast<float4 *>(Cg + (size_t)(row0 + rr) * p.ldc + nbase + c);
                            if (p.accumulate) {
                                const float4 cc = *dst;
                                q.x += cc.x; q.y += cc.y; q.z += cc.z; q.w += cc.w;
                            }
                            *dst = q;
                        }
                    }
                    __syncwarp();
                } else if (row_ok) {
                    float *dst = p.C32 + c_off + nbase;
#pragma unroll
                    for (int e = 0; e < 32; ++e) if (nbase + e < p.N) dst[e] = p.accumulate ? dst[e] + o[e] : o[e];
                }
            }
            if (row_ok) {
                if (p.C16) {
                    __nv_bfloat16 *dst = p.C16 + c_off + nbase;
                    if (nbase + 32 <= p.N && (p.ldc & 7) == 0 && (((uintptr_t)dst & 15u) == 0)) {
#pragma unroll
                        for (int e = 0; e < 32; e += 8)
                            *reinterpret_cast<uint4 *>(dst + e) = make_uint4(pack_bf16x2(o[e], o[e + 1]), pack_bf16x2(o[e + 2], o[e + 3]),
                                                                            pack_bf16x2(o[e + 4], o[e + 5]), pack_bf16x2(o[e + 6], o[e + 7]));
                    } else {
#pragma unroll
                        for (int e = 0; e < 32; ++e) if (nbase + e < p.N) dst[e] = __float2bfloat16_rn(o[e]);
                    }
                }
                if (p.Csoftmax && c0 == 0) {      // N <= 32: the whole row is in this chunk
                    float mx = -3.0e38f;
#pragma unroll
                    for (int e = 0; e < 32; ++e) if (e < p.N) mx = fmaxf(mx, o[e]);
                    float sum = 0.f;
#pragma unroll
                    for (int e = 0; e < 32; ++e) { o[e] = (e < p.N) ? __expf(o[e] - mx) : 0.f; sum += o[e]; }
                    const float inv = 1.0f / sum;
                    float *dst = p.Csoftmax + c_off;
#pragma unroll
                    for (int e = 0; e < 32; ++e) if (e < p.N) dst[e] = o[e] * inv;
                }
            }
        }
    } else if (warp == 4) {
        // ---------------- MMA issuer ----------------
        const uint32_t idesc = make_idesc_bf16(128, (uint32_t)bnt);
        const uint32_t lo_base = (smem_u32(base) & 0x3FFFFu) >> 4;
        for (int kt = 0; kt < KT; ++kt) {
            const int s = kt % kGemmStages;
            mbar_wait(&full[s], (uint32_t)((kt / kGemmStages) & 1), 42);
            tc_fence_after();
            const uint32_t a_lo = lo_base + (uint32_t)s * (kGemmStageBytes >> 4), b_lo = a_lo + (16384u >> 4);
            if (elect_one_sync()) {
#pragma unroll
                for (int k = 0; k < 4; ++k) mma_bf16_ss(tmem, sdesc_from_lo(a_lo + k * 2), sdesc_from_lo(b_lo + k * 2), idesc, (kt | k) > 0);
                mma_commit(&empty[s]);
                if (kt == KT - 1) mma_commit(&acc_full);
            }
            __syncwarp();
        }
    } else {
        // ---------------- weight-tile loader ----------------
        if (lane == 0) {
            const size_t img_tile_bytes = (size_t)img_rows * 128u;
            const uint8_t *img = p.Bimg + (size_t)g * p.b_group_stride + ((size_t)ntile * (p.K >> 6) + kt0) * img_tile_bytes + (size_t)nsub * b_tile_bytes;
            for (int kt = 0; kt < KT; ++kt) {
                const int s = kt % kGemmStages;
                if (kt >= kGemmStages) mbar_wait(&empty[s], (uint32_t)(((kt / kGemmStages) - 1) & 1), 43);
                mbar_arrive_expect_tx(&full[s], b_tile_bytes);
                bulk_g2s(base + s * kGemmStageBytes + 16384, img + (size_t)kt * img_tile_bytes, b_tile_bytes, &full[s]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 128);
}

static inline uint16_t f2bf_host(float f)
{
    uint32_t u; memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

size_t gemm_b_image_bytes(int K, int N)
{
    const int bnt = tc_gemm_bn_tile(N);
    const int ntiles = (N + bnt - 1) / bnt;
    return (size_t)ntiles * (K / 64) * bnt * 128;
}

void pack_gemm_b_image(const float *W, int K, int N, std::vector<uint16_t> &img)
{
    const int bnt = tc_gemm_bn_tile(N), ntiles = (N + bnt - 1) / bnt, KT = K / 64;
    img.assign(gemm_b_image_bytes(K, N) / 2, 0);
    for (int nt = 0; nt < ntiles; ++nt)
        for (int kt = 0; kt < KT; ++kt) {
            uint16_t *tile = img.data() + ((size_t)nt * KT + kt) * (bnt * 64);
            for (int r = 0; r < bnt; ++r) {
                const int n = nt * bnt + r;
                if (n >= N) continue;
                for (int k = 0; k < 64; ++k) tile[sw128_offset((uint32_t)bnt, (uint32_t)r, (uint32_t)k) / 2] = f2bf_host(W[(size_t)(kt * 64 + k) * N + n]);
            }
        }
}

static inline float bf2f_host(uint16_t h) { uint32_t u = (uint32_t)h << 16; float f; memcpy(&f, &u, 4); return f; }

void pack_gemm_b_image_split3(const float *W, int K, int N, std::vector<uint16_t> &img)
{
    std::vector<float> W3((size_t)3 * K * N);
    for (size_t i = 0; i < (size_t)K * N; ++i) {
        const float hi = bf2f_host(f2bf_host(W[i]));
        W3[i] = hi;                                   // rows [0, K): hi(W)    x hi(A)
        W3[(size_t)K * N + i] = W[i] - hi;            // rows [K, 2K): lo(W)   x hi(A)   (rounded to bf16 by the packer)
        W3[(size_t)2 * K * N + i] = hi;               // rows [2K, 3K): hi(W)  x lo(A)
    }
    pack_gemm_b_image(W3.data(), 3 * K, N, img);
}

bool tc_gemm_persist_try(const TcGemmParams &p, cudaStream_t st, int *rc);   // tc_gemm_tma.cu

int tc_gemm(const TcGemmParams &p, cudaStream_t st)
{
    PCP_REQUIRE(p.A && p.Bimg && (p.C32 || p.C16 || p.Csoftmax), PCP_ERR_INVALID, "tc_gemm: null operand");
    PCP_REQUIRE(p.K >= 64 && p.K % 64 == 0, PCP_ERR_UNSUPPORTED, "tc_gemm: K=%d must be a multiple of 64", p.K);
    PCP_REQUIRE(p.N >= 1 && (p.N <= 128 || p.N % 128 == 0), PCP_ERR_UNSUPPORTED, "tc_gemm: N=%d must be <= 128 or a multiple of 128", p.N);
    PCP_REQUIRE(!p.Csoftmax || p.N <= 32, PCP_ERR_UNSUPPORTED, "tc_gemm: softmax epilogue needs N <= 32");
    PCP_REQUIRE((p.lda % (p.a_f32 ? 4 : 8)) == 0, PCP_ERR_UNSUPPORTED, "tc_gemm: lda must keep 16-byte row chunks aligned");
    PCP_REQUIRE(!p.a_split3 || (p.a_f32 && p.K % 192 == 0), PCP_ERR_UNSUPPORTED, "tc_gemm: a_split3 needs fp32 activations and K = 3 x a multiple of 64");
    PCP_REQUIRE(p.ksplit >= 1 && (p.K / 64) % p.ksplit == 0, PCP_ERR_UNSUPPORTED, "tc_gemm: K / 64 = %d is not a multiple of ksplit = %d", p.K / 64, p.ksplit);
    PCP_REQUIRE(p.ksplit == 1 || (p.C32 && !p.C16 && !p.Csoftmax && !p.bias && !p.relu && !p.a_split3), PCP_ERR_INVALID,
                "tc_gemm: split-K adds plain fp32 products");
    if (p.M <= 0 || p.groups <= 0) return PCP_OK;
    if (p.ksplit == 1) {
        int rc = PCP_OK;
        if (tc_gemm_persist_try(p, st, &rc)) return rc;
    }
    const int bnt_img = tc_gemm_bn_tile(p.N);   // MMA N (multiple of 16); the epilogue reads 32-column TMEM chunks and masks n >= N
    const int ntiles = (p.N + bnt_img - 1) / bnt_img;
    const int mtiles = (p.M + 127) / 128;
    const int64_t gy = (int64_t)mtiles * p.groups;
    static bool attr = false;
    if (!attr) {
        PCP_CUDA_OK(cudaFuncSetAttribute(tc_gemm_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem));
        PCP_CUDA_OK(cudaFuncSetAttribute(tc_gemm_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem));
        PCP_CUDA_OK(cudaFuncSetAttribute(tc_gemm_kernel<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem));
        attr = true;
    }
    PCP_REQUIRE(gy <= 65535, PCP_ERR_UNSUPPORTED, "tc_gemm: %lld row tiles exceed one launch", (long long)gy);
    static const bool no_pdl = getenv("PCP_TC_GEMM_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {};
    // few CTAs and full-width tiles: 32-column blocks (see the kernel)
    static const bool no_split = getenv("PCP_TC_GEMM_NO_NSPLIT") != nullptr;
    const int nsplit = (!no_split && bnt_img == 128 && (int64_t)ntiles * gy * 4 <= 2 * (int64_t)sm_count()) ? 4 : 1;
    // the ring only uses min(K / 64, stages) stages
    const int stages_used = std::min(p.K / 64 / p.ksplit, kGemmStages);
    static const bool no_dense = getenv("PCP_TC_GEMM_NO_DENSE") != nullptr;
    const bool dense = !no_dense && p.ksplit == 1 && stages_used <= 2 && (int64_t)ntiles * nsplit * gy >= 3 * (int64_t)sm_count();
    cfg.gridDim = dim3(ntiles * nsplit, (unsigned)gy, (unsigned)p.ksplit); cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = (size_t)stages_used * kGemmStageBytes + 1024; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = no_pdl ? 0 : 1;
    if (p.ksplit > 1) PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, tc_gemm_kernel<2, true>, p, bnt_img / nsplit, mtiles, bnt_img, nsplit));
    else if (dense) PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, tc_gemm_kernel<3, false>, p, bnt_img / nsplit, mtiles, bnt_img, nsplit));
    else PCP_CUDA_OK(cudaLaunchKernelEx(&cfg, tc_gemm_kernel<2, false>, p, bnt_img / nsplit, mtiles, bnt_img, nsplit));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

}  // namespace pcp

// test hook: C[g] (m x n) = act(bf16(A[g]) (m x k) . bf16(W) (k x n) + bias[g]) through tc_gemm with bf16 activations in HBM
// (large shapes take the persistent TMA kernel, small ones the per-tile kernel).  out_kind 0 fp32, 1 bf16, 2 softmax (n <= 32);
// host in / host out (h_c always fp32), synchronous.  h_bias is (groups x n) or null.
extern "C" int pcp_selftest_tc_gemm_bf16(const float *h_a, const float *h_w, const float *h_bias, int64_t m, int groups, int k, int n,
                                         int out_kind, int relu, float *h_c)
{
    using namespace pcp;
    PCP_REQUIRE(h_a && h_w && h_c && m >= 1 && groups >= 1 && k >= 64 && n >= 1, PCP_ERR_INVALID, "pcp_selftest_tc_gemm_bf16: bad arguments");
    const size_t rows = (size_t)m * groups;
    std::vector<uint16_t> a16(rows * k), img;
    for (size_t i = 0; i < a16.size(); ++i) a16[i] = f2bf_host(h_a[i]);
    pack_gemm_b_image(h_w, k, n, img);
    uint16_t *da = nullptr; uint8_t *dimg = nullptr; float *dbias = nullptr; void *dc = nullptr;
    const size_t c_bytes = rows * n * (out_kind == 1 ? 2 : 4);
    PCP_CUDA_OK(cudaMalloc(&da, a16.size() * 2)); PCP_CUDA_OK(cudaMalloc(&dimg, img.size() * 2)); PCP_CUDA_OK(cudaMalloc(&dc, c_bytes));
    PCP_CUDA_OK(cudaMemcpy(da, a16.data(), a16.size() * 2, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(dimg, img.data(), img.size() * 2, cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemset(dc, 0xff, c_bytes));
    if (h_bias) {
        PCP_CUDA_OK(cudaMalloc(&dbias, (size_t)groups * n * 4));
        PCP_CUDA_OK(cudaMemcpy(dbias, h_bias, (size_t)groups * n * 4, cudaMemcpyHostToDevice));
    }
    TcGemmParams p;
    p.A = da; p.a_f32 = 0; p.lda = k; p.a_group_stride = (int64_t)m * k; p.Bimg = dimg; p.bias = dbias; p.bias_group_stride = n;
    p.ldc = n; p.c_group_stride = (int64_t)m * n; p.M = (int)m; p.N = n; p.K = k; p.groups = groups; p.relu = relu;
    if (out_kind == 0) p.C32 = static_cast<float *>(dc);
    else if (out_kind == 1) p.C16 = static_cast<__nv_bfloat16 *>(dc);
    else p.Csoftmax = static_cast<float *>(dc);
    int rc = tc_gemm(p, nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    if (rc == PCP_OK && e != cudaSuccess) { set_error("pcp_selftest_tc_gemm_bf16: %s", cudaGetErrorString(e)); rc = PCP_ERR_CUDA; }
    if (rc == PCP_OK) {
        if (out_kind == 1) {
            std::vector<uint16_t> c16(rows * n);
            cudaMemcpy(c16.data(), dc, c_bytes, cudaMemcpyDeviceToHost);
            for (size_t i = 0; i < c16.size(); ++i) { const uint32_t u = (uint32_t)c16[i] << 16; memcpy(&h_c[i], &u, 4); }
        } else {
            cudaMemcpy(h_c, dc, c_bytes, cudaMemcpyDeviceToHost);
        }
    }
    cudaFree(da); cudaFree(dimg); cudaFree(dc); cudaFree(dbias);
    return rc;
}
