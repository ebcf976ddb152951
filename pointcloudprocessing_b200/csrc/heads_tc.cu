// heads_tc.cu — the small-M dense layers and the segmentation head of the bf16 precision mode on tensor cores
// (tc_gemm.cu):  T-Net tails 1024->512->256->K^2 (pointnet/PointNet.py:431-442), classification head
// 1024->512->256->C + softmax (:252-263), segmentation head 1088->512->256->128->128->P + softmax (:268-290).
//
// Segmentation head, per point, without materialising tile/concat or X_64:
//   h0 = relu(L0(pc_n @ R))                  CUDA-core kernel (K = 3), bf16 (M,64)
//   h1 = relu(h0 . W[s1_l2_64])              tc_gemm
//   s1 = relu(h1 . (R64_b . W1[0:64]) + g_b . W1[64:1088] + t1)    per-cloud folded weight image + per-cloud bias
//   s2..s4, logits, softmax                  tc_gemm chain; activations stay bf16 in HBM between layers
#include <cmath>
#include <cstring>
#include <vector>

#include "model.cuh"
#include "tail_fused.cuh"
#include "tc_gemm.cuh"
#include "umma.cuh"

namespace pcp {

using namespace umma;

struct TcLayerImg {
    uint8_t *img = nullptr;
    const uint4 *frag = nullptr;         // the same weights in mma.sync fragment order (tail_fused.cu), dense tails only
    const float *bias = nullptr;
    int K = 0, N = 0;
};

struct TcHeads {
    TcLayerImg it_d1, it_d2, it_w, ft_d1, ft_d2, ft_w, cls1, cls2, cls3;
    TcLayerImg ft_fold;                  // 256 -> 4096: feature T-Net output folded into s2_l1_64, columns in image order
    TcLayerImg s1_2, seg1g, seg2, seg3, seg4, seg5;
    float *l0_s1 = nullptr;              // [64][4] folded s1_l1_64 (w0,w1,w2,t)
    const float *wseg1a = nullptr;       // fp32 folded W1[0:64, :] (64 x 512) inside d_fold
    const float *seg1_t = nullptr;
    // workspace
    __nv_bfloat16 *d512 = nullptr, *d256 = nullptr; int cap_b = 0;
    __nv_bfloat16 *h0 = nullptr, *h1 = nullptr, *s512 = nullptr, *s256 = nullptr, *s128a = nullptr, *s128b = nullptr;
    uint8_t *wseg_img = nullptr;         // per-cloud folded (R64 . W1[0:64]) images, 64 KB each
    float *gvec = nullptr;
    int64_t cap_pts = 0; int cap_clouds = 0;
    std::vector<void *> allocs;
};

static int upload(TcHeads *h, const void *src, size_t bytes, void **dst)
{
    PCP_CUDA_OK(cudaMalloc(dst, bytes));
    h->allocs.push_back(*dst);
    PCP_CUDA_OK(cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice));
    return PCP_OK;
}

static int make_img(TcHeads *h, TcLayerImg &L, const float *W, const float *d_bias, int K, int N, bool tail = false)
{
    std::vector<uint16_t> img;
    pack_gemm_b_image(W, K, N, img);
    PCP_TRY(upload(h, img.data(), img.size() * 2, reinterpret_cast<void **>(&L.img)));
    if (tail) {          // dense tails also run fused at small batch
        std::vector<uint32_t> fr;
        pack_tail_weights(W, K, N, fr);
        void *d = nullptr;
        PCP_TRY(upload(h, fr.data(), fr.size() * 4, &d));
        L.frag = static_cast<const uint4 *>(d);
    }
    L.bias = d_bias; L.K = K; L.N = N;
    return PCP_OK;
}

// the three dense layers behind a pooled feature in one launch (b <= 32); PCP_NO_FUSED_TAIL keeps the three tc_gemm launches
static bool fused_tail_ok(int b, const TcLayerImg &l1, const TcLayerImg &l2, const TcLayerImg &l3)
{
    static const bool off = getenv("PCP_NO_FUSED_TAIL") != nullptr;
    return !off && b <= kTailMaxRows && l1.frag && l2.frag && l3.frag && l1.K == 1024 && l1.N == 512 && l2.K == 512 && l2.N == 256 && l3.K == 256 &&
           (l3.N <= 32 || l3.N % 2 == 0);
}
static int fused_tail(const float *pooled, int b, const TcLayerImg &l1, const TcLayerImg &l2, const TcLayerImg &l3, float *out32, __nv_bfloat16 *out16,
                      float *out_softmax, cudaStream_t st)
{
    TailFusedParams p;
    p.A0 = pooled; p.b = b; p.w1 = l1.frag; p.w2 = l2.frag; p.w3 = l3.frag; p.t1 = l1.bias; p.t2 = l2.bias; p.t3 = l3.bias; p.n3 = l3.N;
    p.out32 = out32; p.out16 = out16; p.out_softmax = out_softmax;
    return tail_fused(p, st);
}

void heads_tc_destroy(pcp_model *m)
{
    if (!m->tc_heads) return;
    TcHeads *h = reinterpret_cast<TcHeads *>(m->tc_heads);
    for (void *p : h->allocs) PCP_RELEASE(cudaFree(p));
    PCP_RELEASE(cudaFree(h->d512)); PCP_RELEASE(cudaFree(h->d256));
    PCP_RELEASE(cudaFree(h->h0)); PCP_RELEASE(cudaFree(h->h1)); PCP_RELEASE(cudaFree(h->s512)); PCP_RELEASE(cudaFree(h->s256)); PCP_RELEASE(cudaFree(h->s128a)); PCP_RELEASE(cudaFree(h->s128b));
    PCP_RELEASE(cudaFree(h->wseg_img)); PCP_RELEASE(cudaFree(h->gvec));
    delete h;
    m->tc_heads = nullptr;
}

int heads_tc_prepare(pcp_model *m, const std::vector<float> &fold)
{
    heads_tc_destroy(m);
    TcHeads *h = new TcHeads();
    m->tc_heads = h;
    auto layer = [&](TcLayerImg &L, int id, bool tail = false) {
        const LayerSpec &S = m->layers[id];
        return make_img(h, L, fold.data() + S.fold_w, m->d_fold + S.fold_t, S.cin, S.cout, tail);
    };
    if (!m->desc.vanilla) {
        PCP_TRY(layer(h->it_d1, L_IT_D1, true)); PCP_TRY(layer(h->it_d2, L_IT_D2, true));
        PCP_TRY(make_img(h, h->it_w, fold.data() + m->fold_it_w, m->d_fold + m->fold_it_t, 256, 9, true));
        PCP_TRY(layer(h->ft_d1, L_FT_D1, true)); PCP_TRY(layer(h->ft_d2, L_FT_D2, true));
        PCP_TRY(make_img(h, h->ft_w, fold.data() + m->fold_ft_w, m->d_fold + m->fold_ft_t, 256, 4096, true));
        // X_64 = X @ R64 only ever feeds s2_l1_64 in the trunk (PointNet.py:228,236), and R64 = reshape(h . w + b) is linear in
        // the T-Net's last hidden vector h, so the per-cloud operand image (R64 . W)[n, kk] the third chain kernel reads is
        // ONE small-M GEMM of h against T[k, (n, kk)] = sum_m w[k, kk*64 + m] W[m, n] with its columns laid out in the image's
        // SWIZZLE_128B order — no R64 tensor, no per-cloud fold kernel (which was 6.6 % of the forward).
        {
            const float *wt = fold.data() + m->fold_ft_w, *bt = fold.data() + m->fold_ft_t, *W = fold.data() + m->layers[L_S2_1].fold_w;
            std::vector<float> T((size_t)256 * 4096), tb(4096);
            std::vector<uint32_t> pos(4096);
            for (int kk = 0; kk < 64; ++kk)
                for (int n = 0; n < 64; ++n) pos[kk * 64 + n] = sw128_offset(64, (uint32_t)n, (uint32_t)kk) / 2;
            float row[64];
            for (int k = 0; k <= 256; ++k) {                              // k == 256: the bias row
                const float *src = k < 256 ? wt + (size_t)k * 4096 : bt;
                float *dst = k < 256 ? T.data() + (size_t)k * 4096 : tb.data();
                for (int kk = 0; kk < 64; ++kk) {
                    for (int n = 0; n < 64; ++n) row[n] = 0.f;
                    for (int mm = 0; mm < 64; ++mm) {
                        const float r = src[kk * 64 + mm];
                        const float *wr = W + mm * 64;
                        for (int n = 0; n < 64; ++n) row[n] = fmaf(r, wr[n], row[n]);
                    }
                    for (int n = 0; n < 64; ++n) dst[pos[kk * 64 + n]] = row[n];
                }
            }
            float *d_tb = nullptr;
            PCP_TRY(upload(h, tb.data(), tb.size() * 4, reinterpret_cast<void **>(&d_tb)));
            PCP_TRY(make_img(h, h->ft_fold, T.data(), d_tb, 256, 4096, true));
        }
    }
    PCP_TRY(layer(h->cls1, L_CLS1, true)); PCP_TRY(layer(h->cls2, L_CLS2, true)); PCP_TRY(layer(h->cls3, L_CLS3, true));
    PCP_TRY(layer(h->s1_2, L_S1_2));
    const LayerSpec &S1 = m->layers[L_SEG1];
    // global half of seg_l1_512: rows 64..1087 of the folded kernel (1024 x 512)
    PCP_TRY(make_img(h, h->seg1g, fold.data() + S1.fold_w + (size_t)64 * 512, m->d_fold + S1.fold_t, 1024, 512));
    h->wseg1a = m->d_fold + S1.fold_w;
    h->seg1_t = m->d_fold + S1.fold_t;
    PCP_TRY(layer(h->seg2, L_SEG2)); PCP_TRY(layer(h->seg3, L_SEG3)); PCP_TRY(layer(h->seg4, L_SEG4)); PCP_TRY(layer(h->seg5, L_SEG5));
    const LayerSpec &L0 = m->layers[L_S1_1];
    std::vector<float> l0(256);
    for (int ch = 0; ch < 64; ++ch) {
        for (int k = 0; k < 3; ++k) l0[ch * 4 + k] = fold[L0.fold_w + (size_t)k * 64 + ch];
        l0[ch * 4 + 3] = fold[L0.fold_t + ch];
    }
    PCP_TRY(upload(h, l0.data(), l0.size() * 4, reinterpret_cast<void **>(&h->l0_s1)));
    return PCP_OK;
}

static int gemm_layer(const TcLayerImg &L, const void *A, int a_f32, int64_t lda, int M, int relu, float *C32, __nv_bfloat16 *C16,
                      float *Csoftmax, int64_t ldc, cudaStream_t st)
{
    TcGemmParams p;
    p.A = A; p.a_f32 = a_f32; p.lda = lda; p.Bimg = L.img; p.bias = L.bias;
    p.C32 = C32; p.C16 = C16; p.Csoftmax = Csoftmax; p.ldc = ldc; p.M = M; p.N = L.N; p.K = L.K; p.relu = relu;
    return tc_gemm(p, st);
}

static int ensure_small(TcHeads *h, int b)
{
    if (h->cap_b >= b) return PCP_OK;
    cudaFree(h->d512); cudaFree(h->d256);
    h->d512 = h->d256 = nullptr; h->cap_b = 0;
    PCP_CUDA_OK(cudaMalloc(&h->d512, (size_t)b * 512 * 2));
    PCP_CUDA_OK(cudaMalloc(&h->d256, (size_t)b * 256 * 2));
    h->cap_b = b;
    return PCP_OK;
}

// TNet tail (PointNet.py:431-442): pooled (b,1024) fp32 -> 512 -> 256 -> K^2 (+ b), fp32 out
int heads_tc_tnet_tail(pcp_model *m, int which, const float *pooled, int b, float *out, cudaStream_t st)
{
    TcHeads *h = reinterpret_cast<TcHeads *>(m->tc_heads);
    PCP_TRY(ensure_small(h, b));
    const TcLayerImg &d1 = which == 0 ? h->it_d1 : h->ft_d1, &d2 = which == 0 ? h->it_d2 : h->ft_d2, &w = which == 0 ? h->it_w : h->ft_w;
    if (fused_tail_ok(b, d1, d2, w)) return fused_tail(pooled, b, d1, d2, w, out, nullptr, nullptr, st);
    PCP_TRY(gemm_layer(d1, pooled, 1, 1024, b, 1, nullptr, h->d512, nullptr, 512, st));
    PCP_TRY(gemm_layer(d2, h->d512, 0, 512, b, 1, nullptr, h->d256, nullptr, 256, st));
    return gemm_layer(w, h->d256, 0, 256, b, 0, out, nullptr, nullptr, w.N, st);
}

// feature T-Net tail straight into the per-cloud (R64 . W[s2_l1_64]) operand images of the third chain kernel (b x 8 KB)
// (r64_out != null: also the plain 64 x 64 matrices, which the segmentation head folds into its own first layer)
int heads_tc_tnet_tail_fold(pcp_model *m, const float *pooled, int b, uint8_t *wc_img, float *r64_out, cudaStream_t st)
{
    TcHeads *h = reinterpret_cast<TcHeads *>(m->tc_heads);
    PCP_TRY(ensure_small(h, b));
    if (!r64_out && fused_tail_ok(b, h->ft_d1, h->ft_d2, h->ft_fold))
        return fused_tail(pooled, b, h->ft_d1, h->ft_d2, h->ft_fold, nullptr, reinterpret_cast<__nv_bfloat16 *>(wc_img), nullptr, st);
    PCP_TRY(gemm_layer(h->ft_d1, pooled, 1, 1024, b, 1, nullptr, h->d512, nullptr, 512, st));
    PCP_TRY(gemm_layer(h->ft_d2, h->d512, 0, 512, b, 1, nullptr, h->d256, nullptr, 256, st));
    PCP_TRY(gemm_layer(h->ft_fold, h->d256, 0, 256, b, 0, nullptr, reinterpret_cast<__nv_bfloat16 *>(wc_img), nullptr, 4096, st));
    if (r64_out) PCP_TRY(gemm_layer(h->ft_w, h->d256, 0, 256, b, 0, r64_out, nullptr, nullptr, 4096, st));
    return PCP_OK;
}

// classification head (PointNet.py:252-263); Dropout is the identity at inference
int heads_tc_cls(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st)
{
    TcHeads *h = reinterpret_cast<TcHeads *>(m->tc_heads);
    PCP_TRY(ensure_small(h, b));
    const int C = m->desc.num_classes;
    if (C <= 32 && fused_tail_ok(b, h->cls1, h->cls2, h->cls3)) return fused_tail(g3, b, h->cls1, h->cls2, h->cls3, d_cls_logits, nullptr, d_cls, st);
    PCP_TRY(gemm_layer(h->cls1, g3, 1, 1024, b, 1, nullptr, h->d512, nullptr, 512, st));
    PCP_TRY(gemm_layer(h->cls2, h->d512, 0, 512, b, 1, nullptr, h->d256, nullptr, 256, st));
    if (C <= 32)
        return gemm_layer(h->cls3, h->d256, 0, 256, b, 0, d_cls_logits, nullptr, d_cls, C, st);
    float *logits = d_cls_logits ? d_cls_logits : m->ws.clslog;
    PCP_TRY(gemm_layer(h->cls3, h->d256, 0, 256, b, 0, logits, nullptr, nullptr, C, st));
    if (d_cls) PCP_TRY(softmax_rows(logits, d_cls, b, C, st));
    return PCP_OK;
}

// h0 = relu(L0((x - c)/s [@ R])) as bf16 (M,64): the K = 3 layer on CUDA cores, 128 B/point written
__global__ void __launch_bounds__(256)
seg_l0_kernel(const float *__restrict__ x, int64_t n, int64_t npts, const float *__restrict__ centroid, const float *__restrict__ scale,
              const float *__restrict__ R, const float *__restrict__ l0, __nv_bfloat16 *__restrict__ h0)
{
    __shared__ alignas(16) float w[256];
    w[threadIdx.x] = l0[threadIdx.x];
    __syncthreads();
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < npts; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t cloud = i / n;
        const float sc = __ldg(scale + cloud);
        float q0 = __fdiv_rn(__fsub_rn(__ldg(x + 3 * i), __ldg(centroid + cloud * 3)), sc);
        float q1 = __fdiv_rn(__fsub_rn(__ldg(x + 3 * i + 1), __ldg(centroid + cloud * 3 + 1)), sc);
        float q2 = __fdiv_rn(__fsub_rn(__ldg(x + 3 * i + 2), __ldg(centroid + cloud * 3 + 2)), sc);
        if (R) {
            const float *r = R + cloud * 9;
            const float a = q0 * __ldg(r) + q1 * __ldg(r + 3) + q2 * __ldg(r + 6);
            const float b = q0 * __ldg(r + 1) + q1 * __ldg(r + 4) + q2 * __ldg(r + 7);
            const float c = q0 * __ldg(r + 2) + q1 * __ldg(r + 5) + q2 * __ldg(r + 8);
            q0 = a; q1 = b; q2 = c;
        }
        uint4 *dst = reinterpret_cast<uint4 *>(h0 + i * 64);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            float o[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const float4 ww = *reinterpret_cast<const float4 *>(&w[(c * 8 + e) * 4]);
                const float v = fmaf(q2, ww.z, fmaf(q1, ww.y, fmaf(q0, ww.x, ww.w)));
                o[e] = v > 0.f ? v : 0.f;
            }
            dst[c] = make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
        }
    }
}

// per-cloud fold for the local half of seg_l1_512: image_b(n, k) = bf16( sum_m R64_b[k][m] * W[m][n] ), n < 512, k < 64;
// written as the tc_gemm weight image (4 tiles of 128 rows x 64 K).  R64 == nullptr (vanilla): R64 = I.
__global__ void __launch_bounds__(256)
fold_seg1_kernel(const float *__restrict__ R64, const float *__restrict__ W, uint8_t *__restrict__ img)
{
    __shared__ __align__(16) float sRt[64 * 68];            // R64 transposed: sRt[m][k], rows padded to 68 floats (16-byte aligned)
    __shared__ __align__(16) float sW[64 * 64];
    const int64_t cloud = blockIdx.y;
    const int c0 = blockIdx.x * 64;               // 64-column slice of the 512 outputs
    {
        // all 32 loads of a thread are issued before the first shared store: one memory round trip, not sixteen
        float rr[16], ww[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int i = threadIdx.x + j * 256;
            rr[j] = R64 ? __ldg(R64 + cloud * 4096 + i) : ((i >> 6) == (i & 63) ? 1.f : 0.f);
            ww[j] = __ldg(W + (size_t)(i >> 6) * 512 + c0 + (i & 63));
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int i = threadIdx.x + j * 256;
            sRt[(i & 63) * 68 + (i >> 6)] = rr[j];
            sW[i] = ww[j];
        }
    }
    __syncthreads();
    uint8_t *out = img + cloud * 65536 + (size_t)(c0 >> 7) * 16384;
    // thread -> (output column n, 16 consecutive k): the image row of column n holds its 64 k-values contiguously (128 B), so a
    // thread's results leave as two 16-byte stores (the first version scattered 2-byte stores: 73 us for 128 clouds)
    const int nl = threadIdx.x & 63, k0 = (threadIdx.x >> 6) * 16;
    float acc[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = 0.f;
#pragma unroll 4
    for (int mm = 0; mm < 64; ++mm) {
        const float w = sW[mm * 64 + nl];
        const float4 *r4 = reinterpret_cast<const float4 *>(&sRt[mm * 68 + k0]);       // warp-uniform address: broadcast
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 r = r4[q];
            acc[4 * q] = fmaf(r.x, w, acc[4 * q]); acc[4 * q + 1] = fmaf(r.y, w, acc[4 * q + 1]);
            acc[4 * q + 2] = fmaf(r.z, w, acc[4 * q + 2]); acc[4 * q + 3] = fmaf(r.w, w, acc[4 * q + 3]);
        }
    }
    const uint32_t row = (uint32_t)((c0 & 127) + nl);
#pragma unroll
    for (int h8 = 0; h8 < 2; ++h8)
        *reinterpret_cast<uint4 *>(out + sw128_offset(128, row, (uint32_t)(k0 + 8 * h8))) =
            make_uint4(pack_bf16x2(acc[8 * h8], acc[8 * h8 + 1]), pack_bf16x2(acc[8 * h8 + 2], acc[8 * h8 + 3]),
                       pack_bf16x2(acc[8 * h8 + 4], acc[8 * h8 + 5]), pack_bf16x2(acc[8 * h8 + 6], acc[8 * h8 + 7]));
}

static int ensure_seg(TcHeads *h, int clouds, int64_t pts)
{
    if (h->cap_pts >= pts && h->cap_clouds >= clouds) return PCP_OK;
    cudaFree(h->h0); cudaFree(h->h1); cudaFree(h->s512); cudaFree(h->s256); cudaFree(h->s128a); cudaFree(h->s128b);
    cudaFree(h->wseg_img); cudaFree(h->gvec);
    h->h0 = h->h1 = h->s512 = h->s256 = h->s128a = h->s128b = nullptr; h->wseg_img = nullptr; h->gvec = nullptr;
    h->cap_pts = 0; h->cap_clouds = 0;
    PCP_CUDA_OK(cudaMalloc(&h->h0, (size_t)pts * 64 * 2)); PCP_CUDA_OK(cudaMalloc(&h->h1, (size_t)pts * 64 * 2));
    PCP_CUDA_OK(cudaMalloc(&h->s512, (size_t)pts * 512 * 2)); PCP_CUDA_OK(cudaMalloc(&h->s256, (size_t)pts * 256 * 2));
    PCP_CUDA_OK(cudaMalloc(&h->s128a, (size_t)pts * 128 * 2)); PCP_CUDA_OK(cudaMalloc(&h->s128b, (size_t)pts * 128 * 2));
    PCP_CUDA_OK(cudaMalloc(&h->wseg_img, (size_t)clouds * 65536));
    PCP_CUDA_OK(cudaMalloc(&h->gvec, (size_t)clouds * 512 * 4));
    h->cap_pts = pts; h->cap_clouds = clouds;
    return PCP_OK;
}

int heads_tc_seg(pcp_model *m, const float *d_x, int b, int64_t n, const float *centroid, const float *scale, const float *d_R,
                 const float *d_R64, const float *g3, float *d_seg, float *d_seg_logits, cudaStream_t st)
{
    TcHeads *h = reinterpret_cast<TcHeads *>(m->tc_heads);
    const int P = m->desc.num_parts;
    PCP_REQUIRE(P <= 32, PCP_ERR_UNSUPPORTED, "bf16 segmentation head supports up to 32 part labels (got %d)", P);
    const int64_t max_pts = (int64_t)1 << 19;                  // 512k points per chunk: ~1.2 GB of bf16 activations
    const int cb = (int)std::max<int64_t>(1, std::min<int64_t>(b, max_pts / n));
    PCP_TRY(ensure_seg(h, cb, (int64_t)cb * n));
    for (int c0 = 0; c0 < b; c0 += cb) {
        const int bc = std::min(cb, b - c0);
        const int64_t M = (int64_t)bc * n;
        PCP_REQUIRE(M < ((int64_t)1 << 31), PCP_ERR_UNSUPPORTED, "segmentation chunk too large");
        const float *x = d_x + (int64_t)c0 * n * 3;
        int grid = (int)std::min<int64_t>((M + 255) / 256, (int64_t)sm_count() * 8);
        seg_l0_kernel<<<grid, 256, 0, st>>>(x, n, M, centroid + (int64_t)c0 * 3, scale + c0, d_R ? d_R + (int64_t)c0 * 9 : nullptr, h->l0_s1, h->h0);
        PCP_CHECK_LAUNCH();
        PCP_TRY(gemm_layer(h->s1_2, h->h0, 0, 64, (int)M, 1, nullptr, h->h1, nullptr, 64, st));               // :220
        fold_seg1_kernel<<<dim3(8, bc), 256, 0, st>>>(d_R64 ? d_R64 + (int64_t)c0 * 4096 : nullptr, h->wseg1a, h->wseg_img);
        PCP_CHECK_LAUNCH();
        // per-cloud bias: global . W1[64:1088] + t1  (the tile/concat of PointNet.py:268-270 never exists)
        PCP_TRY(gemm_layer(h->seg1g, g3 + (int64_t)c0 * 1024, 1, 1024, bc, 0, h->gvec, nullptr, nullptr, 512, st));
        TcGemmParams p;
        p.A = h->h1; p.lda = 64; p.a_group_stride = n * 64; p.Bimg = h->wseg_img; p.b_group_stride = 65536;
        p.bias = h->gvec; p.bias_group_stride = 512; p.C16 = h->s512; p.ldc = 512; p.c_group_stride = n * 512;
        p.M = (int)n; p.N = 512; p.K = 64; p.groups = bc; p.relu = 1;
        PCP_TRY(tc_gemm(p, st));                                                                              // :275
        PCP_TRY(gemm_layer(h->seg2, h->s512, 0, 512, (int)M, 1, nullptr, h->s256, nullptr, 256, st));         // :278
        PCP_TRY(gemm_layer(h->seg3, h->s256, 0, 256, (int)M, 1, nullptr, h->s128a, nullptr, 128, st));        // :281
        PCP_TRY(gemm_layer(h->seg4, h->s128a, 0, 128, (int)M, 1, nullptr, h->s128b, nullptr, 128, st));       // :284
        PCP_TRY(gemm_layer(h->seg5, h->s128b, 0, 128, (int)M, 0, d_seg_logits ? d_seg_logits + (int64_t)c0 * n * P : nullptr, nullptr,
                           d_seg ? d_seg + (int64_t)c0 * n * P : nullptr, P, st));                            // :287 + softmax
    }
    return PCP_OK;
}

}  // namespace pcp
