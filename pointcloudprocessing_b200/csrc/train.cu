// train.cu — the training step of pointnet_train.py:310-351 on the device: training-mode forward of
// PointNet.call (pointnet/PointNet.py:197-292: batch-statistics BatchNormalization on thawed layers,
// Dropout), the three losses (SCCE on probabilities with the 1e-7 clip, MSE on R, T-Net regulariser
// PointNet.py:447-451), the full backward pass into one flat gradient buffer, and fused keras Adam.
//
// Structure: every weighted layer is Z = A . W (GEMM), then BN (batch or moving statistics) + ReLU as an
// elementwise pass.  Pre-BN activations Z are kept for the backward pass, and so are the trunk's post-activation
// tensors (TL::Y); the heads' are recomputed from Z when a weight gradient needs them.  The reduce_max over points
// (PointNet.py:248,429) is fused with BN + ReLU into one column-max pass and its gradient is split evenly between
// tied maxima (the tf.reduce_max convention) — fp32 mode; the bf16 mode handles the pooled layers
// without their (rows x 1024) tensors (train_pool.inl).  GEMMs go through mm_nn / mm_nt / mm_tn below: fp32 CUDA-core
// tiles in PCP_PRECISION_FP32, bf16 tcgen05 tiles with fp32 accumulation in PCP_PRECISION_BF16 (tc_gemm.cu,
// tc_gemm_tma.cu, tc_train.cu), their weight images re-packed once per step (ImgCache).
//
// Per-column reductions (BN statistics, BN backward sums, the 3-channel weight gradients) end in per-block partial
// sums that one block per element adds in block order (reduce_parts_kernel / bn_finalize_kernel): no same-address
// atomics, the same bits on every run.  train_step_impl splits the step into a prelude (allocation, tables), the
// stream-ordered body (train_step_body) and the metric read-back; the body is replayed as one CUDA graph once a
// configuration has run twice.
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "gemm_simt.cuh"
#include "model.cuh"
#include "tc_gemm.cuh"
#include "tc_train.cuh"

namespace pcp {

static const float kBnEpsT = 1e-3f;       // keras BatchNormalization epsilon
static const float kBnMomentum = 0.99f;   // PointNet.py:502,528

// ------------------------------------------------------------------------------------------------
// Fused keras Adam over the flat parameter buffer (pointnet_train.py:317-319):
//   m = b1*m + (1-b1)*g ; v = b2*v + (1-b2)*g^2 ; w -= alpha * m / (sqrt(v) + eps),
//   alpha = lr * sqrt(1-b2^t) / (1-b1^t).  28 B/parameter of HBM traffic (read w,g,m,v; write w,m,v).
// `mask` holds one byte per 4-float group: 0 => frozen / BN statistic, skipped.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
adam_kernel(float4 *__restrict__ w, const float4 *__restrict__ g, float4 *__restrict__ m, float4 *__restrict__ v,
            const unsigned char *__restrict__ mask, size_t n4, float alpha, float b1, float b2, float eps)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        if (!mask[i]) continue;
        float4 W = w[i], G = g[i], M = m[i], V = v[i];
        float *pw = &W.x, *pg = &G.x, *pm = &M.x, *pv = &V.x;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            pm[j] = b1 * pm[j] + (1.f - b1) * pg[j];
            pv[j] = b2 * pv[j] + (1.f - b2) * pg[j] * pg[j];
            pw[j] = pw[j] - alpha * pm[j] / (sqrtf(pv[j]) + eps);
        }
        w[i] = W; m[i] = M; v[i] = V;
    }
}

// ------------------------------------------------------------------------------------------------
// Column-wise kernels over a row-major (rows x C) activation.  C is a multiple of 64, C <= 1024.
// Thread mapping (256 threads): C >= 256 -> each thread owns C/256 strided columns of one row lane;
// C < 256 -> 256/C row lanes run side by side.  Loads are coalesced across the columns of a row.
// ------------------------------------------------------------------------------------------------
struct BnP {
    const float *mean, *sc, *beta, *inv;      // pre = (z - mean) * sc + beta ; xhat = (z - mean) * inv ; sc = gamma * inv
};

__device__ __forceinline__ float bn_pre(float z, float mean, float sc, float beta) { return fmaf(z - mean, sc, beta); }

struct ColMap {
    int cpt, lanes, col0, lane;
    __device__ ColMap(int C) {
        cpt = C >= 256 ? C / 256 : 1;
        lanes = C >= 256 ? 1 : 256 / C;
        col0 = C >= 256 ? (int)threadIdx.x : (int)threadIdx.x % C;
        lane = C >= 256 ? 0 : (int)threadIdx.x / C;
    }
};

// cross-lane reduction of up to 4 per-thread doubles into lane 0 (shared scratch: 256 doubles)
__device__ __forceinline__ void lane_reduce(double *sh, const ColMap &cm, int C, double (&v)[4])
{
    if (cm.lanes == 1) return;
    __syncthreads();
    sh[threadIdx.x] = v[0];
    __syncthreads();
    if (cm.lane == 0) {
        double s = 0;
        for (int l = 0; l < cm.lanes; ++l) s += sh[l * C + cm.col0];
        v[0] = s;
    }
}

// sum / sum of squares per column (BatchNormalization batch statistics, biased variance)
__global__ void __launch_bounds__(256)
colstats_kernel(const float *__restrict__ Z, int64_t rows, int C, double *__restrict__ parts)
{
    double *sum = parts + (size_t)blockIdx.x * 2 * C, *sumsq = sum + C;
    __shared__ double sh[256];
    const ColMap cm(C);
    const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(rows, r0 + per);
    double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
    for (int64_t r = r0 + cm.lane; r < r1; r += cm.lanes) {
        const float *zr = Z + r * C;
#pragma unroll 4
        for (int j = 0; j < cm.cpt; ++j) {
            const double z = (double)__ldg(zr + cm.col0 + j * 256);
            s[j] += z; q[j] += z * z;
        }
    }
    lane_reduce(sh, cm, C, s);
    lane_reduce(sh, cm, C, q);
    if (cm.lane == 0)
        for (int j = 0; j < cm.cpt; ++j) {
            sum[cm.col0 + j * 256] = s[j];
            sumsq[cm.col0 + j * 256] = q[j];
        }
}

// float4 variants of the two column reductions (every BN layer of the network: C in {64 ... 1024}).  A thread owns four
// adjacent columns and strides over the rows with four independent 16-byte loads in flight (the scalar versions above
// left one 4-byte load per thread outstanding: ~1 TB/s); fp64 accumulation throughout, like the scalar versions.
struct ColMap4 {
    int c4, lanes, col, lane;          // C/4 float4 columns; 256 / c4 row lanes (c4 <= 256)
    __device__ ColMap4(int C) { c4 = C >> 2; lanes = 256 / c4; col = (int)threadIdx.x % c4; lane = (int)threadIdx.x / c4; }
};
static inline bool colmap4_ok(int C) { return C % 4 == 0 && C / 4 <= 256 && 256 % (C / 4) == 0; }

__device__ __forceinline__ void lane_reduce4(double (*sh)[256], const ColMap4 &cm, double (&v)[4])
{
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 4; ++e) sh[e][threadIdx.x] = v[e];
    __syncthreads();
    if (cm.lane == 0)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            double s = 0;
            for (int l = 0; l < cm.lanes; ++l) s += sh[e][l * cm.c4 + cm.col];
            v[e] = s;
        }
}

__global__ void __launch_bounds__(256)
colstats4_kernel(const float4 *__restrict__ Z, int64_t rows, int C, double *__restrict__ parts)
{
    double *sum = parts + (size_t)blockIdx.x * 2 * C, *sumsq = sum + C;
    __shared__ double sh[4][256];
    const ColMap4 cm(C);
    const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(rows, r0 + per);
    double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
    auto acc = [&](const float4 z) {
        const double x = z.x, y = z.y, w = z.z, v = z.w;
        s[0] += x; s[1] += y; s[2] += w; s[3] += v;
        q[0] = fma(x, x, q[0]); q[1] = fma(y, y, q[1]); q[2] = fma(w, w, q[2]); q[3] = fma(v, v, q[3]);
    };
    int64_t r = r0 + cm.lane;
    const int64_t step = cm.lanes;
    for (; r + 3 * step < r1; r += 4 * step) {
        const float4 z0 = __ldg(Z + r * cm.c4 + cm.col), z1 = __ldg(Z + (r + step) * cm.c4 + cm.col),
                     z2 = __ldg(Z + (r + 2 * step) * cm.c4 + cm.col), z3 = __ldg(Z + (r + 3 * step) * cm.c4 + cm.col);
        acc(z0); acc(z1); acc(z2); acc(z3);
    }
    for (; r < r1; r += step) acc(__ldg(Z + r * cm.c4 + cm.col));
    lane_reduce4(sh, cm, s);
    lane_reduce4(sh, cm, q);
    if (cm.lane == 0)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            sum[cm.col * 4 + e] = s[e];
            sumsq[cm.col * 4 + e] = q[e];
        }
}

__global__ void __launch_bounds__(256)
bn_bwd_reduce4_kernel(const float4 *__restrict__ dY, const float4 *__restrict__ Z, int64_t rows, int C, BnP p, int relu,
                      double *__restrict__ parts)
{
    double *s1 = parts + (size_t)blockIdx.x * 2 * C, *s2 = s1 + C;
    __shared__ double sh[4][256];
    const ColMap4 cm(C);
    const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(rows, r0 + per);
    const float4 mu = *reinterpret_cast<const float4 *>(p.mean + cm.col * 4), sc = *reinterpret_cast<const float4 *>(p.sc + cm.col * 4),
                 be = *reinterpret_cast<const float4 *>(p.beta + cm.col * 4), iv = *reinterpret_cast<const float4 *>(p.inv + cm.col * 4);
    double a[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
    auto acc = [&](const float4 z, float4 dy) {
        const float zx = z.x - mu.x, zy = z.y - mu.y, zz = z.z - mu.z, zw = z.w - mu.w;
        if (relu) {
            if (!(fmaf(zx, sc.x, be.x) > 0.f)) dy.x = 0.f;
            if (!(fmaf(zy, sc.y, be.y) > 0.f)) dy.y = 0.f;
            if (!(fmaf(zz, sc.z, be.z) > 0.f)) dy.z = 0.f;
            if (!(fmaf(zw, sc.w, be.w) > 0.f)) dy.w = 0.f;
        }
        a[0] += (double)dy.x; a[1] += (double)dy.y; a[2] += (double)dy.z; a[3] += (double)dy.w;
        b[0] = fma((double)dy.x, (double)(zx * iv.x), b[0]); b[1] = fma((double)dy.y, (double)(zy * iv.y), b[1]);
        b[2] = fma((double)dy.z, (double)(zz * iv.z), b[2]); b[3] = fma((double)dy.w, (double)(zw * iv.w), b[3]);
    };
    int64_t r = r0 + cm.lane;
    const int64_t step = cm.lanes;
    for (; r + step < r1; r += 2 * step) {
        const int64_t i0 = r * cm.c4 + cm.col, i1 = (r + step) * cm.c4 + cm.col;
        const float4 z0 = __ldg(Z + i0), d0 = __ldg(dY + i0), z1 = __ldg(Z + i1), d1 = __ldg(dY + i1);
        acc(z0, d0); acc(z1, d1);
    }
    for (; r < r1; r += step) acc(__ldg(Z + r * cm.c4 + cm.col), __ldg(dY + r * cm.c4 + cm.col));
    lane_reduce4(sh, cm, a);
    lane_reduce4(sh, cm, b);
    if (cm.lane == 0)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            s1[cm.col * 4 + e] = a[e];
            s2[cm.col * 4 + e] = b[e];
        }
}

// dW (3 x C) += X^T (3 x rows) . dZ (rows x C): the two 3-channel input layers (a 128 x 64 GEMM tile wastes 125 of its rows)
__global__ void __launch_bounds__(256)
wgrad3_kernel(const float *__restrict__ X, const float4 *__restrict__ dZ, int64_t rows, int C, double *__restrict__ parts)
{
    double *dW = parts + (size_t)blockIdx.x * 3 * C;
    __shared__ double sh[4][256];
    const ColMap4 cm(C);
    const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(rows, r0 + per);
    float4 a0 = make_float4(0.f, 0.f, 0.f, 0.f), a1 = a0, a2 = a0;
    auto acc3 = [&](const float4 d, float x, float y, float z) {
        a0.x = fmaf(x, d.x, a0.x); a0.y = fmaf(x, d.y, a0.y); a0.z = fmaf(x, d.z, a0.z); a0.w = fmaf(x, d.w, a0.w);
        a1.x = fmaf(y, d.x, a1.x); a1.y = fmaf(y, d.y, a1.y); a1.z = fmaf(y, d.z, a1.z); a1.w = fmaf(y, d.w, a1.w);
        a2.x = fmaf(z, d.x, a2.x); a2.y = fmaf(z, d.y, a2.y); a2.z = fmaf(z, d.z, a2.z); a2.w = fmaf(z, d.w, a2.w);
    };
    int64_t r = r0 + cm.lane;
    const int64_t step = cm.lanes;
    for (; r + 3 * step < r1; r += 4 * step) {                    // four rows in flight
        float4 d[4]; float x[4], y[4], z[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t rr = r + u * step;
            d[u] = __ldg(dZ + rr * cm.c4 + cm.col); x[u] = __ldg(X + rr * 3); y[u] = __ldg(X + rr * 3 + 1); z[u] = __ldg(X + rr * 3 + 2);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) acc3(d[u], x[u], y[u], z[u]);
    }
    for (; r < r1; r += step) acc3(__ldg(dZ + r * cm.c4 + cm.col), __ldg(X + r * 3), __ldg(X + r * 3 + 1), __ldg(X + r * 3 + 2));
    const float4 acc[3] = {a0, a1, a2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double v[4] = {acc[k].x, acc[k].y, acc[k].z, acc[k].w};
        lane_reduce4(sh, cm, v);
        if (cm.lane == 0)
#pragma unroll
            for (int e = 0; e < 4; ++e) dW[(size_t)k * C + cm.col * 4 + e] = v[e];
    }
}

// The column reductions above end in per-block partial sums (parts[block][...]) instead of one atomic per column and
// block: 592 blocks' double atomics on the same 128 words serialise in the L2 at ~40 ns each, a fixed 25 us per launch
// that was most of colstats4 (36 us for 67 MB) and bn_bwd_reduce4.  The partials are summed in block order — the same
// bits on every run — by one warp per element here, or inside bn_finalize_kernel.
// sum_p col[p * stride] over the block's 256 threads: every thread's (<= 4) loads leave together — ONE L2 round trip for
// 592 partials where a warp per element walked five dependent ones (7 us) — and the partial sums are added in a fixed
// order (lanes by shuffle, then the eight warps in turn).  Result valid in thread 0.
__device__ __forceinline__ double block_parts_sum(const double *__restrict__ col, size_t stride, int nparts, double *sh8)
{
    const int tid = threadIdx.x;
    double v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) { const int p = tid + 256 * u; v[u] = p < nparts ? col[(size_t)p * stride] : 0.0; }
    double s = (v[0] + v[1]) + (v[2] + v[3]);
    for (int p = tid + 1024; p < nparts; p += 256) s += col[(size_t)p * stride];
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __syncthreads();
    if ((tid & 31) == 0) sh8[tid >> 5] = s;
    __syncthreads();
    return ((sh8[0] + sh8[1]) + (sh8[2] + sh8[3])) + ((sh8[4] + sh8[5]) + (sh8[6] + sh8[7]));
}

//   out[i] = sum_p parts[p][i];  f_lo / f_hi: float copies of the two halves (dbeta / dgamma);  add_f32[i] += sum.  Block = element.
__global__ void __launch_bounds__(256)
reduce_parts_kernel(const double *__restrict__ parts, int nparts, int count, double *__restrict__ out, float *__restrict__ f_lo,
                    float *__restrict__ f_hi, int half, float *__restrict__ add_f32)
{
    __shared__ double sh8[8];
    const int i = blockIdx.x;
    const double s = block_parts_sum(parts + i, (size_t)count, nparts, sh8);
    if (threadIdx.x) return;
    if (out) out[i] = s;
    if (f_lo && i < half) f_lo[i] = (float)s;
    if (f_hi && i >= half) f_hi[i - half] = (float)s;
    if (add_f32) add_f32[i] += (float)s;
}

// batch statistics -> (mean, inv, sc, beta) and the moving-average update; or the moving statistics when the
// layer is frozen (keras: training and trainable).  One block per column; parts[p][0:C] are the partial
// sums, parts[p][C:2C] the partial sums of squares (nparts = 1: already reduced)
__global__ void __launch_bounds__(256)
bn_finalize_kernel(const double *__restrict__ parts, int nparts, double rows, int C, int batch, const float *__restrict__ gamma,
                   const float *__restrict__ beta_in, float *__restrict__ mov_mean, float *__restrict__ mov_var, float *__restrict__ mean,
                   float *__restrict__ inv, float *__restrict__ sc, float *__restrict__ beta, float eps, float momentum)
{
    __shared__ double sh8[2][8];
    const int c = blockIdx.x;
    double sum = 0, sumsq = 0;
    if (batch) {                                       // both sums in one pass: their loads leave together, one pair of barriers
        const int tid = threadIdx.x;
        const size_t stride = (size_t)2 * C;
        double a[4], q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int p = tid + 256 * u;
            a[u] = p < nparts ? parts[(size_t)p * stride + c] : 0.0;
            q[u] = p < nparts ? parts[(size_t)p * stride + C + c] : 0.0;
        }
        double s = (a[0] + a[1]) + (a[2] + a[3]), s2 = (q[0] + q[1]) + (q[2] + q[3]);
        for (int p = tid + 1024; p < nparts; p += 256) { s += parts[(size_t)p * stride + c]; s2 += parts[(size_t)p * stride + C + c]; }
#pragma unroll
        for (int o = 16; o; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
        if ((tid & 31) == 0) { sh8[0][tid >> 5] = s; sh8[1][tid >> 5] = s2; }
        __syncthreads();
        sum = ((sh8[0][0] + sh8[0][1]) + (sh8[0][2] + sh8[0][3])) + ((sh8[0][4] + sh8[0][5]) + (sh8[0][6] + sh8[0][7]));
        sumsq = ((sh8[1][0] + sh8[1][1]) + (sh8[1][2] + sh8[1][3])) + ((sh8[1][4] + sh8[1][5]) + (sh8[1][6] + sh8[1][7]));
    }
    if (threadIdx.x) return;
    float mu, var;
    if (batch) {
        const double m = sum / rows;
        double v = sumsq / rows - m * m;
        if (v < 0) v = 0;
        mu = (float)m; var = (float)v;
        mov_mean[c] = momentum * mov_mean[c] + (1.f - momentum) * mu;
        mov_var[c] = momentum * mov_var[c] + (1.f - momentum) * var;
    } else {
        mu = mov_mean[c]; var = mov_var[c];
    }
    const float iv = 1.0f / sqrtf(var + eps);
    mean[c] = mu; inv[c] = iv; sc[c] = gamma[c] * iv; beta[c] = beta_in[c];
}

// Y = relu((Z - mean) * sc + beta)
__global__ void __launch_bounds__(256)
bn_act_kernel(const float4 *__restrict__ Z, int64_t n4, int C4, BnP p, float4 *__restrict__ Y, uint2 *__restrict__ Y16)
{   // Y16 (optional): the same activation rounded to bf16 for the GEMM of the next layer (persistent TMA kernel)
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4) * 4;
        const float4 z = __ldg(Z + i);
        const float4 mu = __ldg(reinterpret_cast<const float4 *>(p.mean + c)), s = __ldg(reinterpret_cast<const float4 *>(p.sc + c)),
                     be = __ldg(reinterpret_cast<const float4 *>(p.beta + c));
        float4 y;
        y.x = fmaxf(bn_pre(z.x, mu.x, s.x, be.x), 0.f); y.y = fmaxf(bn_pre(z.y, mu.y, s.y, be.y), 0.f);
        y.z = fmaxf(bn_pre(z.z, mu.z, s.z, be.z), 0.f); y.w = fmaxf(bn_pre(z.w, mu.w, s.w, be.w), 0.f);
        Y[i] = y;
        if (Y16) {
            const __nv_bfloat162 lo = __floats2bfloat162_rn(y.x, y.y), hi = __floats2bfloat162_rn(y.z, y.w);
            Y16[i] = make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
        }
    }
}

// per cloud: gmax[cloud, c] = max over the cloud's points of relu(bn(Z))   (zero-initialised; values >= 0)
__global__ void __launch_bounds__(256)
bn_relu_colmax_kernel(const float *__restrict__ Z, int64_t n, int C, BnP p, float *__restrict__ gmax)
{
    const ColMap cm(C);
    const int64_t cloud = blockIdx.y;
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(n, r0 + per);
    float mu[4], s[4], be[4], mx[4];
    for (int j = 0; j < cm.cpt; ++j) {
        const int c = cm.col0 + j * 256;
        mu[j] = p.mean[c]; s[j] = p.sc[c]; be[j] = p.beta[c]; mx[j] = 0.f;
    }
    for (int64_t r = r0 + cm.lane; r < r1; r += cm.lanes) {
        const float *zr = Z + (cloud * n + r) * C;
#pragma unroll 4
        for (int j = 0; j < cm.cpt; ++j) mx[j] = fmaxf(mx[j], bn_pre(__ldg(zr + cm.col0 + j * 256), mu[j], s[j], be[j]));
    }
    if (r1 > r0)
        for (int j = 0; j < cm.cpt; ++j)
            atomicMax(reinterpret_cast<unsigned int *>(gmax + cloud * C + cm.col0 + j * 256), __float_as_uint(mx[j]));
}

// tie count of the per-cloud maximum (even gradient split of tf.reduce_max)
__global__ void __launch_bounds__(256)
pool_count_kernel(const float *__restrict__ Z, int64_t n, int C, BnP p, const float *__restrict__ gmax, int *__restrict__ cnt)
{
    const ColMap cm(C);
    const int64_t cloud = blockIdx.y;
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(n, r0 + per);
    float mu[4], s[4], be[4], g[4];
    int k[4] = {0, 0, 0, 0};
    for (int j = 0; j < cm.cpt; ++j) {
        const int c = cm.col0 + j * 256;
        mu[j] = p.mean[c]; s[j] = p.sc[c]; be[j] = p.beta[c]; g[j] = gmax[cloud * C + c];
    }
    for (int64_t r = r0 + cm.lane; r < r1; r += cm.lanes) {
        const float *zr = Z + (cloud * n + r) * C;
#pragma unroll 4
        for (int j = 0; j < cm.cpt; ++j) {
            const float y = fmaxf(bn_pre(__ldg(zr + cm.col0 + j * 256), mu[j], s[j], be[j]), 0.f);
            k[j] += (y == g[j]) ? 1 : 0;
        }
    }
    if (r1 > r0)
        for (int j = 0; j < cm.cpt; ++j)
            if (k[j]) atomicAdd(cnt + cloud * C + cm.col0 + j * 256, k[j]);
}

// dY[p, c] = dG[cloud, c] / cnt[cloud, c] where relu(bn(Z))[p, c] equals the pooled maximum, else 0
__global__ void __launch_bounds__(256)
pool_bwd_kernel(const float4 *__restrict__ Z, int64_t n4_per_cloud, int64_t total4, int C4, BnP p, const float *__restrict__ gmax,
                const float *__restrict__ dG, const int *__restrict__ cnt, float4 *__restrict__ dY)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total4; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4) * 4;
        const int64_t cloud = i / n4_per_cloud;
        const float4 z = __ldg(Z + i);
        const float zz[4] = {z.x, z.y, z.z, z.w};
        float o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float y = fmaxf(bn_pre(zz[j], p.mean[c + j], p.sc[c + j], p.beta[c + j]), 0.f);
            const int64_t gi = cloud * (C4 * 4) + c + j;
            o[j] = (y == gmax[gi]) ? dG[gi] / (float)cnt[gi] : 0.f;
        }
        dY[i] = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// BN + ReLU backward, reduction half: s1[c] = sum dyh, s2[c] = sum dyh * xhat, dyh = dY * [pre > 0]
__global__ void __launch_bounds__(256)
bn_bwd_reduce_kernel(const float *__restrict__ dY, const float *__restrict__ Z, int64_t rows, int C, BnP p, int relu,
                     double *__restrict__ parts)
{
    double *s1 = parts + (size_t)blockIdx.x * 2 * C, *s2 = s1 + C;
    __shared__ double sh[256];
    const ColMap cm(C);
    const int64_t per = (rows + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(rows, r0 + per);
    float mu[4], s[4], be[4], iv[4];
    double a[4] = {0, 0, 0, 0}, b[4] = {0, 0, 0, 0};
    for (int j = 0; j < cm.cpt; ++j) {
        const int c = cm.col0 + j * 256;
        mu[j] = p.mean[c]; s[j] = p.sc[c]; be[j] = p.beta[c]; iv[j] = p.inv[c];
    }
    for (int64_t r = r0 + cm.lane; r < r1; r += cm.lanes) {
#pragma unroll 4
        for (int j = 0; j < cm.cpt; ++j) {
            const int64_t idx = r * C + cm.col0 + j * 256;
            const float zc = __ldg(Z + idx) - mu[j];
            float dy = __ldg(dY + idx);
            if (relu && !(fmaf(zc, s[j], be[j]) > 0.f)) dy = 0.f;
            a[j] += (double)dy; b[j] += (double)dy * (double)(zc * iv[j]);
        }
    }
    lane_reduce(sh, cm, C, a);
    lane_reduce(sh, cm, C, b);
    if (cm.lane == 0)
        for (int j = 0; j < cm.cpt; ++j) {
            s1[cm.col0 + j * 256] = a[j];
            s2[cm.col0 + j * 256] = b[j];
        }
}

// BN + ReLU backward, elementwise half (in place on dY):
//   batch statistics : dZ = sc * (dyh - s1/rows - xhat * s2/rows)
//   moving statistics: dZ = sc * dyh
__global__ void __launch_bounds__(256)
bn_bwd_apply_kernel(float4 *__restrict__ dY, const float4 *__restrict__ Z, int64_t n4, int C4, BnP p, int relu, int batch,
                    const double *__restrict__ s1, const double *__restrict__ s2, float inv_rows)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % C4) * 4;
        const float4 z = __ldg(Z + i);
        float4 d = dY[i];
        const float zz[4] = {z.x, z.y, z.z, z.w};
        float dd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float zc = zz[j] - p.mean[c + j];
            float dy = dd[j];
            if (relu && !(fmaf(zc, p.sc[c + j], p.beta[c + j]) > 0.f)) dy = 0.f;
            if (batch) dy = dy - (float)s1[c + j] * inv_rows - (zc * p.inv[c + j]) * ((float)s2[c + j] * inv_rows);
            dd[j] = p.sc[c + j] * dy;
        }
        dY[i] = make_float4(dd[0], dd[1], dd[2], dd[3]);
    }
}

// the same with a thread bound to four adjacent columns (ColMap4): the eight per-column constants live in registers instead
// of being re-read (24 scalar loads and 8 double -> float conversions per float4 pair before: 3.4 TB/s; the loop is now two
// 16-byte loads and one store per element group, two rows in flight)
__global__ void __launch_bounds__(256)
bn_bwd_apply4_kernel(float4 *__restrict__ dY, const float4 *__restrict__ Z, int64_t rows, int C, BnP p, int relu, int batch,
                     const double *__restrict__ s1, const double *__restrict__ s2, float inv_rows, uint2 *__restrict__ dZ16)
{   // dZ16 (optional): the same dZ rounded to bf16 — the input-gradient GEMM that follows then streams it through TMA
    const ColMap4 cm(C);
    const int c = cm.col * 4;
    float mu[4], sc[4], be[4], iv[4], k1[4], k2[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        mu[j] = p.mean[c + j]; sc[j] = p.sc[c + j]; be[j] = p.beta[c + j]; iv[j] = p.inv[c + j];
        k1[j] = batch ? (float)s1[c + j] * inv_rows : 0.f;
        k2[j] = batch ? (float)s2[c + j] * inv_rows : 0.f;
    }
    auto apply = [&](const float4 z, const float4 d) {
        const float zz[4] = {z.x, z.y, z.z, z.w};
        float dd[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float zc = zz[j] - mu[j];
            float dy = dd[j];
            if (relu && !(fmaf(zc, sc[j], be[j]) > 0.f)) dy = 0.f;
            if (batch) dy = dy - k1[j] - (zc * iv[j]) * k2[j];
            dd[j] = sc[j] * dy;
        }
        return make_float4(dd[0], dd[1], dd[2], dd[3]);
    };
    auto put = [&](int64_t i, const float4 v) {
        dY[i] = v;
        if (dZ16) {
            const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            dZ16[i] = make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
        }
    };
    const int64_t L = (int64_t)gridDim.x * cm.lanes;
    int64_t r = (int64_t)blockIdx.x * cm.lanes + cm.lane;
    for (; r + L < rows; r += 2 * L) {
        const int64_t i0 = r * cm.c4 + cm.col, i1 = (r + L) * cm.c4 + cm.col;
        const float4 z0 = __ldg(Z + i0), d0 = dY[i0], z1 = __ldg(Z + i1), d1 = dY[i1];
        put(i0, apply(z0, d0));
        put(i1, apply(z1, d1));
    }
    for (; r < rows; r += L) {
        const int64_t i0 = r * cm.c4 + cm.col;
        put(i0, apply(__ldg(Z + i0), dY[i0]));
    }
}

__global__ void store_bn_grads_kernel(const double *__restrict__ s1, const double *__restrict__ s2, int C, float *__restrict__ dgamma,
                                      float *__restrict__ dbeta)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c < C) { dgamma[c] = (float)s2[c]; dbeta[c] = (float)s1[c]; }
}

// out[g, c] (+)= sum over the group's rows of X[(g*rpg + r), c]   (bias gradients; per-cloud sums of the seg head)
__global__ void __launch_bounds__(256)
group_colsum_kernel(const float *__restrict__ X, int64_t rpg, int C, float *__restrict__ out)
{
    const int64_t g = blockIdx.y;
    const int64_t per = (rpg + gridDim.x - 1) / gridDim.x;
    const int64_t r0 = (int64_t)blockIdx.x * per, r1 = min(rpg, r0 + per);
    for (int c = blockIdx.z * blockDim.x + threadIdx.x; c < C; c += gridDim.z * blockDim.x) {
        float s = 0.f;
        for (int64_t r = r0; r < r1; ++r) s += __ldg(X + (g * rpg + r) * C + c);
        if (r1 > r0) atomicAdd(out + g * C + c, s);
    }
}

// Philox keep-mask (Dropout, PointNet.py:255,260): counter = (global sample id, unit, layer), kept units scaled by 1/(1-rate)
__device__ __forceinline__ float dropout_factor(uint64_t seed, uint64_t sample, uint32_t unit, uint32_t layer, float rate, float scale)
{
    uint32_t c[4] = {(uint32_t)sample, (uint32_t)(sample >> 32), unit, layer};
    philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
    const float u = (float)(c[0] >> 8) * (1.0f / 16777216.0f);
    return u >= rate ? scale : 0.f;
}

// H = Y * factor ; explicit mask (0/1 floats) when given, else Philox.  Also used for the backward (dY = dH * factor).
__global__ void __launch_bounds__(256)
dropout_kernel(const float *Y, int64_t rows, int C, const float *__restrict__ mask, uint64_t seed, int64_t sample0,
               uint32_t layer, float rate, float *H, const unsigned long long *__restrict__ dyn)
{   // dyn (optional, device): {seed, sample0} of THIS step — a captured step graph is replayed with new values every step
    if (dyn) { seed = dyn[0]; sample0 = (int64_t)dyn[1]; }
    const float scale = 1.0f / (1.0f - rate);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < rows * C; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / C;
        const int c = (int)(i % C);
        const float f = mask ? mask[i] * scale : dropout_factor(seed, (uint64_t)(sample0 + r), (uint32_t)c, layer, rate, scale);
        H[i] = Y[i] * f;
    }
}

__global__ void set_step_scalars_kernel(unsigned long long *dyn, unsigned long long seed, long long sample0)
{
    dyn[0] = seed; dyn[1] = (unsigned long long)sample0;
}

// softmax + SCCE on probabilities (keras from_logits=False: clip to [1e-7, 1-1e-7], -log; pointnet_train.py:337-338)
// one thread per row; C <= 64.  dlogits = gscale * (p - onehot) unless the label probability is clipped (zero gradient).
__global__ void __launch_bounds__(256)
softmax_ce_kernel(const float *__restrict__ logits, int64_t rows, int C, const int32_t *__restrict__ labels, float gscale,
                  float *__restrict__ prob, float *__restrict__ dlogits, double *__restrict__ loss_sum, double *__restrict__ correct)
{
    double lsum = 0, csum = 0;
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
        const float *z = logits + r * C;
        float mx = z[0]; int am = 0;
        for (int c = 1; c < C; ++c) if (z[c] > mx) { mx = z[c]; am = c; }
        float s = 0.f;
        for (int c = 0; c < C; ++c) s += expf(z[c] - mx);
        const float inv = 1.0f / s;
        const int y = labels[r];
        const float py = expf(z[y] - mx) * inv;
        const bool clipped = (py < 1e-7f) || (py > 1.0f - 1e-7f);
        const float pc = fminf(fmaxf(py, 1e-7f), 1.0f - 1e-7f);
        lsum += (double)(-logf(pc));
        csum += (am == y) ? 1.0 : 0.0;
        for (int c = 0; c < C; ++c) {
            const float p = expf(z[c] - mx) * inv;
            if (prob) prob[r * C + c] = p;
            if (dlogits) dlogits[r * C + c] = clipped ? 0.f : gscale * (p - (c == y ? 1.f : 0.f));
        }
    }
    lsum = warp_sum(lsum); csum = warp_sum(csum);
    if ((threadIdx.x & 31) == 0) { atomicAdd(loss_sum, lsum); atomicAdd(correct, csum); }
}

// The same per 128-row tile staged through shared memory (C <= 96): the thread-per-row kernel above reads and writes rows
// 4*C bytes apart from neighbouring lanes (113 us for 262 144 x 50, 0.9 TB/s) and ends in one atomic pair per WARP on one
// line; here the tile moves with coalesced 16-byte accesses, the row arithmetic (identical, in the same order) runs on
// shared memory, and a block adds its loss / accuracy sums once.
__global__ void __launch_bounds__(128)
softmax_ce_tile_kernel(const float *__restrict__ logits, int64_t rows, int C, const int32_t *__restrict__ labels, float gscale,
                       float *__restrict__ prob, float *__restrict__ dlogits, double *__restrict__ loss_sum, double *__restrict__ correct)
{
    extern __shared__ __align__(16) float ce_tile[];
    __shared__ double red[2][4];
    double lsum = 0, csum = 0;
    const int64_t ntiles = (rows + 127) / 128;
    const bool vec = ((reinterpret_cast<uintptr_t>(logits) | reinterpret_cast<uintptr_t>(prob) | reinterpret_cast<uintptr_t>(dlogits)) & 15u) == 0;
    auto copy_out = [&](float *dst, int cnt) {
        const int c4 = vec ? cnt >> 2 : 0;
        for (int i = threadIdx.x; i < c4; i += 128) reinterpret_cast<float4 *>(dst)[i] = reinterpret_cast<const float4 *>(ce_tile)[i];
        for (int i = (c4 << 2) + threadIdx.x; i < cnt; i += 128) dst[i] = ce_tile[i];
    };
    for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int64_t r0 = t * 128;
        const int nr = (int)min((int64_t)128, rows - r0), cnt = nr * C;
        const float *src = logits + r0 * C;                     // 128 * C floats per tile: 16-byte aligned with the base
        {
            const int c4 = vec ? cnt >> 2 : 0;
            for (int i = threadIdx.x; i < c4; i += 128) reinterpret_cast<float4 *>(ce_tile)[i] = __ldg(reinterpret_cast<const float4 *>(src) + i);
            for (int i = (c4 << 2) + threadIdx.x; i < cnt; i += 128) ce_tile[i] = __ldg(src + i);
        }
        __syncthreads();
        float *z = ce_tile + threadIdx.x * C;
        bool clipped = false;
        int y = 0;
        if ((int)threadIdx.x < nr) {
            float mx = z[0]; int am = 0;
            for (int c = 1; c < C; ++c) if (z[c] > mx) { mx = z[c]; am = c; }
            float s = 0.f;
            for (int c = 0; c < C; ++c) s += expf(z[c] - mx);
            const float inv = 1.0f / s;
            y = labels[r0 + threadIdx.x];
            const float py = expf(z[y] - mx) * inv;
            clipped = (py < 1e-7f) || (py > 1.0f - 1e-7f);
            const float pc = fminf(fmaxf(py, 1e-7f), 1.0f - 1e-7f);
            lsum += (double)(-logf(pc));
            csum += (am == y) ? 1.0 : 0.0;
            for (int c = 0; c < C; ++c) z[c] = expf(z[c] - mx) * inv;
        }
        __syncthreads();
        if (prob) copy_out(prob + r0 * C, cnt);
        if (dlogits) {
            __syncthreads();
            if ((int)threadIdx.x < nr)
                for (int c = 0; c < C; ++c) z[c] = clipped ? 0.f : gscale * (z[c] - (c == y ? 1.f : 0.f));
            __syncthreads();
            copy_out(dlogits + r0 * C, cnt);
        }
        __syncthreads();
    }
    lsum = warp_sum(lsum); csum = warp_sum(csum);
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = lsum; red[1][threadIdx.x >> 5] = csum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        atomicAdd(loss_sum, (red[0][0] + red[0][1]) + (red[0][2] + red[0][3]));
        atomicAdd(correct, (red[1][0] + red[1][1]) + (red[1][2] + red[1][3]));
    }
}

// Z (rows x C) = X (rows x 3) . W (3 x C) [+ bias]: the two 3-channel input layers (a 128 x 64 x 16 GEMM tile spends 13 of its
// 16 k-steps on padding: 49 us for a 67 MB store).  A thread owns four adjacent columns of one row; fma order of the GEMM.
__global__ void __launch_bounds__(256)
mm3_kernel(const float *__restrict__ X, const float *__restrict__ W, int ldw, const float *__restrict__ bias, int64_t rows, int C4,
           float4 *__restrict__ Z)
{
    const int64_t total = rows * C4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / C4;
        const int c = (int)(i - r * C4) * 4;
        const float x0 = __ldg(X + 3 * r), x1 = __ldg(X + 3 * r + 1), x2 = __ldg(X + 3 * r + 2);
        const float4 w0 = __ldg(reinterpret_cast<const float4 *>(W + c)), w1 = __ldg(reinterpret_cast<const float4 *>(W + ldw + c)),
                     w2 = __ldg(reinterpret_cast<const float4 *>(W + 2 * ldw + c));
        float4 z;
        z.x = fmaf(x2, w2.x, fmaf(x1, w1.x, x0 * w0.x)); z.y = fmaf(x2, w2.y, fmaf(x1, w1.y, x0 * w0.y));
        z.z = fmaf(x2, w2.z, fmaf(x1, w1.z, x0 * w0.z)); z.w = fmaf(x2, w2.w, fmaf(x1, w1.w, x0 * w0.w));
        if (bias) { const float4 b = __ldg(reinterpret_cast<const float4 *>(bias + c)); z.x += b.x; z.y += b.y; z.z += b.z; z.w += b.w; }
        Z[i] = z;
    }
}

// bf16 copy of an fp32 activation (n4 groups of four): lets a GEMM over it take the persistent TMA kernel
__global__ void __launch_bounds__(256)
to_bf16_kernel(const float4 *__restrict__ X, int64_t n4, uint2 *__restrict__ Y)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = __ldg(X + i);
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        Y[i] = make_uint2(*reinterpret_cast<const uint32_t *>(&lo), *reinterpret_cast<const uint32_t *>(&hi));
    }
}

// MSE(R, y_se3) (pointnet_train.py:339): dR += gscale * 2 (R - y); sq_sum += sum (R - y)^2
__global__ void mse_kernel(const float *__restrict__ R, const float *__restrict__ y, int64_t count, float gscale, float *__restrict__ dR,
                           double *__restrict__ sq_sum)
{
    double s = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const float d = R[i] - y[i];
        s += (double)d * d;
        if (dR) dR[i] += gscale * 2.f * d;
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) atomicAdd(sq_sum, s);
}

// T-Net orthogonality regulariser (PointNet.py:447-451): L = 1e-3 * sum_b 1/2 ||I - m m^T||_F^2 ; dm += -2e-3 (I - m m^T) m
// one block per cloud; k = 3 or 64 (shared memory 2 * k*k floats)
__global__ void __launch_bounds__(256)
tnet_reg_kernel(const float *__restrict__ Mx, int k, float *__restrict__ dM, double *__restrict__ loss_sum)
{
    extern __shared__ float shm[];
    float *m = shm, *D = shm + k * k;
    __shared__ double red[8];
    const float *src = Mx + (int64_t)blockIdx.x * k * k;
    for (int i = threadIdx.x; i < k * k; i += blockDim.x) m[i] = src[i];
    __syncthreads();
    double l = 0;
    for (int i = threadIdx.x; i < k * k; i += blockDim.x) {
        const int r = i / k, c = i % k;
        float acc = 0.f;
        for (int j = 0; j < k; ++j) acc = fmaf(m[r * k + j], m[c * k + j], acc);
        const float d = (r == c ? 1.f : 0.f) - acc;
        D[i] = d;
        l += 0.5 * (double)d * d;
    }
    __syncthreads();
    if (dM) {
        float *dst = dM + (int64_t)blockIdx.x * k * k;
        for (int i = threadIdx.x; i < k * k; i += blockDim.x) {
            const int r = i / k, c = i % k;
            float acc = 0.f;
            for (int j = 0; j < k; ++j) acc = fmaf(D[r * k + j], m[j * k + c], acc);
            dst[i] += -2.0e-3f * acc;
        }
    }
    l = warp_sum(l);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = l;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        atomicAdd(loss_sum, 1.0e-3 * t);
    }
}

int tnet_reg_accumulate(const float *M, int b, int k, double *acc, cudaStream_t st)
{
    if (b <= 0) return PCP_OK;
    tnet_reg_kernel<<<b, k <= 8 ? 32 : 256, 2 * (size_t)k * k * sizeof(float), st>>>(M, k, nullptr, acc);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

__global__ void axpy_kernel(float *__restrict__ y, const float *__restrict__ x, int64_t n, float a)
{
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] += a * x[i];
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
// The packed bf16 images of the step's weight matrices.  The weights only change in the Adam step, so every image the
// previous step used is re-packed by ONE launch at the start of the step (41 launches of ~4 us each before); a matrix
// seen for the first time is packed on the spot and joins the table.
struct ImgCache {
    uint8_t *pool = nullptr;
    size_t cap = 0, used = 0;
    std::vector<PackJob> jobs;
    PackJob *d_jobs = nullptr;
    int d_count = 0;              // jobs in the device table
    int fresh = 0;                // jobs[0 .. fresh) were packed by this step's refresh
    const float *lo = nullptr, *hi = nullptr;   // the model's parameter buffer: only matrices inside it are cached (anything else — the
                                                // pooled layers' per-step Q — lives in the arena and changes within the step)
};
// buffers of one implicitly handled max-pooled layer (train_pool.inl)
struct PoolBufs {
    unsigned long long *ext = nullptr;   // 2 x (b x cout) packed extrema keys (max, then min)
    float *gram = nullptr;               // cin x cin : A^T A of the bf16-rounded input (this rank's rows)
    float *gram_parts = nullptr;         // per-split partial Gram matrices (tc_gram_det)
    double *asum = nullptr;              // cin       : 1^T A
    double *colparts = nullptr;          // per-block partial column sums (pool_colsum_kernel), 4 x SMs x 128
    float *gram_g = nullptr;             // the same summed over the ranks (sync-BN); aliases gram / asum otherwise
    double *asum_g = nullptr;
    float *zsel = nullptr;               // b x cout  : the selected extremum of z
    int *arg = nullptr;                  // b x cout  : its row inside the cloud
    float *S = nullptr;                  // b x cout  : sc * dY at that row
    float *c0 = nullptr, *c1 = nullptr;  // cout
    float *Wt = nullptr;                 // cout x cin: W^T, bf16-rounded
    float *Q = nullptr, *q = nullptr;    // cin x cin, cin
};

struct TL {                       // one weighted layer's training record
    int id = -1;
    float *Z = nullptr;           // pre-BN output (rows x cout)
    float *Y = nullptr;           // relu(bn(Z)) kept for the backward pass (trunk layers; 0.7 GB at 256 x 1024 instead of seven recomputations)
    __nv_bfloat16 *Y16 = nullptr; // the same rounded to bf16 (the 128-wide inputs of the pooled layers: their backward GEMM streams it by TMA)
    int64_t rows = 0;
    int cin = 0, cout = 0;
    bool has_bn = false, batch = false, trainable = false;
    float *mean = nullptr, *inv = nullptr, *sc = nullptr, *beta = nullptr;
    double *red = nullptr;        // 2 * cout doubles (forward: sum / sumsq ; backward: s1 / s2)
    BnP bnp() const { return BnP{mean, sc, beta, inv}; }
};

struct TrainState {
    unsigned char *d_mask = nullptr;
    bool mask_valid = false;
    bool trainable_seen[4] = {false, false, false, false};
    // activation arena
    int b = 0; int64_t n = 0;
    char *arena = nullptr; size_t arena_bytes = 0;
    TL L[L_COUNT];
    float *pcn = nullptr, *X3 = nullptr, *X2 = nullptr, *X64 = nullptr;
    float *g1 = nullptr, *g2 = nullptr, *g3 = nullptr;             // pooled (b x 1024)
    int *cnt = nullptr;                                            // (b x 1024) tie counts
    float *R = nullptr, *R64 = nullptr, *dR = nullptr, *dR64 = nullptr;
    float *Ys[8] = {nullptr};                                      // small-layer activations (rows = b)
    float *H1 = nullptr, *H2 = nullptr;                            // after dropout
    float *cls_logits = nullptr, *cls_prob = nullptr, *seg_logits = nullptr, *seg_prob = nullptr;
    float *gvec = nullptr, *dgvec = nullptr;                       // per-cloud bias of seg_l1 (b x 512) and its gradient
    double *colparts = nullptr;                                    // per-block partial sums of the column reductions
    __nv_bfloat16 *G16 = nullptr;                                  // bf16 copy of the dZ the next input-gradient GEMM reads (M x 128)
    float *fshift = nullptr;                                       // 512 floats: epilogue constant of a frozen layer (seg_head_fwd_frozen)
    float *dg1 = nullptr, *dg2 = nullptr, *dg3 = nullptr;
    float *dsmall[4] = {nullptr};                                  // (b x 512), (b x 256), (b x C) ...
    // per-point scratch
    float *Y64a = nullptr, *Y64b = nullptr, *Y128 = nullptr, *Y128b = nullptr, *Y256 = nullptr, *Y512 = nullptr;
    float *G1024 = nullptr, *G512 = nullptr, *G256 = nullptr, *G128a = nullptr, *G128b = nullptr, *G64a = nullptr, *G64b = nullptr;
    float *GP = nullptr, *dX64 = nullptr, *dX2 = nullptr, *dX3 = nullptr;
    double *metrics = nullptr;                                     // 8 doubles
    bool have_seg = false;
    char *eval_buf = nullptr; size_t eval_bytes = 0;               // pcp_model_evaluate: metrics + logits
    uint8_t *tc_img = nullptr;                                     // scratch bf16 weight image of the tensor-core GEMMs (4 MB)
    ImgCache img_cache;                                            // the step's packed weight images (bf16 mode)
    // the captured step (train_step_impl): one graph launch instead of ~220 kernel launches, which is what a small shard
    // (32 clouds per GPU in the 8-GPU strong-scaling run) spends its time on
    struct StepKey {
        int b; long long n; int global_batch; float w_cls, w_seg, w_rot; int skip, tr[4], has_cls, has_seg, has_se3, jobs;
    };
    StepKey graph_key = {}; int graph_seen = 0;                    // consecutive eager steps with this key
    cudaGraphExec_t graph_exec = nullptr;
    cudaStream_t graph_stream = nullptr;                           // the graph's own stream (the caller's may be the legacy default
    cudaEvent_t graph_in = nullptr, graph_out = nullptr;           // stream, which cannot be captured): fork / join by events
    bool graph_broken = false;                                     // a capture failed: stay eager
    float *stage_x = nullptr, *stage_se3 = nullptr; int32_t *stage_ycls = nullptr, *stage_yseg = nullptr;   // the graph's inputs
    unsigned long long *step_scalars = nullptr;                    // {dropout seed, sample offset} of the step being replayed
    void *norm_ws = nullptr; size_t norm_ws_bytes = 0;             // pcp_normalize_ws_f32 scratch (LiDAR-sized clouds)
    PoolBufs pool[3];                                              // L_IT_C3, L_FT_C3, L_S2_3 (bf16 mode)
    bool implicit_pool = false;                                    // pooled layers without their (rows x 1024) tensors (train_pool.inl)
    pcp_allreduce_fn sync_fn = nullptr; void *sync_user = nullptr; // data-parallel BatchNorm (pcp_model_set_bn_allreduce)
};

static int ensure_train_buffers(pcp_model *m)
{
    if (!m->d_grads) {
        PCP_CUDA_OK(cudaMalloc(&m->d_grads, m->n_params * sizeof(float)));
        PCP_CUDA_OK(cudaMalloc(&m->d_adam_m, m->n_params * sizeof(float)));
        PCP_CUDA_OK(cudaMalloc(&m->d_adam_v, m->n_params * sizeof(float)));
        PCP_CUDA_OK(cudaMemset(m->d_grads, 0, m->n_params * sizeof(float)));
        PCP_CUDA_OK(cudaMemset(m->d_adam_m, 0, m->n_params * sizeof(float)));
        PCP_CUDA_OK(cudaMemset(m->d_adam_v, 0, m->n_params * sizeof(float)));
    }
    if (!m->train_ws) m->train_ws = new TrainState();
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    bool same = ts->mask_valid;
    for (int g = 0; g < 4; ++g) same = same && (ts->trainable_seen[g] == m->trainable[g]);
    if (!same) {
        std::vector<unsigned char> mask(m->n_params / 4, 0);
        for (const TensorSpec &t : m->tensors) {
            if (t.is_stat || !m->trainable[t.group]) continue;
            for (size_t i = t.offset / 4; i < (t.offset + t.count + 3) / 4; ++i) mask[i] = 1;
        }
        if (!ts->d_mask) PCP_CUDA_OK(cudaMalloc(&ts->d_mask, mask.size()));
        PCP_CUDA_OK(cudaMemcpy(ts->d_mask, mask.data(), mask.size(), cudaMemcpyHostToDevice));
        for (int g = 0; g < 4; ++g) ts->trainable_seen[g] = m->trainable[g];
        ts->mask_valid = true;
    }
    return PCP_OK;
}

void train_destroy(pcp_model *m)
{
    if (!m->train_ws) return;
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    PCP_RELEASE(cudaFree(ts->d_mask));
    PCP_RELEASE(cudaFree(ts->arena));
    PCP_RELEASE(cudaFree(ts->eval_buf));
    PCP_RELEASE(cudaFree(ts->tc_img));
    PCP_RELEASE(cudaFree(ts->img_cache.pool)); PCP_RELEASE(cudaFree(ts->img_cache.d_jobs));
    if (ts->graph_exec) PCP_RELEASE(cudaGraphExecDestroy(ts->graph_exec));
    if (ts->graph_stream) PCP_RELEASE(cudaStreamDestroy(ts->graph_stream));
    if (ts->graph_in) PCP_RELEASE(cudaEventDestroy(ts->graph_in));
    if (ts->graph_out) PCP_RELEASE(cudaEventDestroy(ts->graph_out));
    delete ts;
    m->train_ws = nullptr;
}

// capacity of the column reductions' partial-sum scratch: blocks (<= 8 per SM) x (2 or 3) x up to 1024 columns
static inline size_t colparts_doubles() { return (size_t)sm_count() * 8 * 3 * 1024; }

// (re)build the activation arena for b clouds of n points
static int ensure_arena(pcp_model *m, TrainState *ts, int b, int64_t n)
{
    if (ts->arena && ts->b == b && ts->n == n) return PCP_OK;
    // bf16 mode handles the pooled 128 -> 1024 layers implicitly; PCP_TRAIN_EXPLICIT_POOL=1 keeps the explicit path (A/B tests)
    static const bool force_explicit = [] { const char *e = getenv("PCP_TRAIN_EXPLICIT_POOL"); return e && e[0] == '1'; }();
    ts->implicit_pool = m->desc.precision == PCP_PRECISION_BF16 && !force_explicit;
    PCP_RELEASE(cudaFree(ts->arena));
    ts->arena = nullptr; ts->b = 0; ts->n = 0;
    const int64_t M = (int64_t)b * n;
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    std::vector<std::pair<void **, size_t>> req;
    auto want = [&](auto **p, size_t bytes) { req.emplace_back(reinterpret_cast<void **>(p), (bytes + 255) & ~(size_t)255); };
    static const int kPointLayers[] = {L_IT_C1, L_IT_C2, L_IT_C3, L_S1_1, L_S1_2, L_FT_C1, L_FT_C2, L_FT_C3, L_S2_1, L_S2_2, L_S2_3,
                                       L_SEG1, L_SEG2, L_SEG3, L_SEG4, L_SEG5};
    for (int i = 0; i < L_COUNT; ++i) {
        TL &t = ts->L[i];
        t = TL();
        const LayerSpec &S = m->layers[i];
        if (!S.present) continue;
        bool per_point = false;
        for (int id : kPointLayers) per_point = per_point || (id == i);
        t.id = i; t.rows = per_point ? M : b; t.cin = S.cin; t.cout = S.cout; t.has_bn = S.has_bn;
        const bool pooled = (i == L_IT_C3 || i == L_FT_C3 || i == L_S2_3);
        if (i != L_SEG5 && i != L_CLS3 && !(pooled && ts->implicit_pool)) want(&t.Z, (size_t)t.rows * S.cout * 4);
        if (i == L_IT_C1 || i == L_IT_C2 || i == L_S1_1 || i == L_FT_C1 || i == L_FT_C2 || i == L_S2_1 || i == L_S2_2)
            want(&t.Y, (size_t)t.rows * S.cout * 4);
        if (i == L_IT_C2 || i == L_FT_C2 || i == L_S2_2) want(&t.Y16, (size_t)t.rows * S.cout * 2);
        if (S.has_bn) {
            want(&t.mean, S.cout * 4); want(&t.inv, S.cout * 4); want(&t.sc, S.cout * 4); want(&t.beta, S.cout * 4);
            want(&t.red, (size_t)S.cout * 2 * 8);
        }
    }
    want(&ts->pcn, M * 12); want(&ts->X3, M * 12); want(&ts->X2, M * 256); want(&ts->X64, M * 256);
    want(&ts->g1, (size_t)b * 4096); want(&ts->g2, (size_t)b * 4096); want(&ts->g3, (size_t)b * 4096); want(&ts->cnt, (size_t)b * 4096);
    want(&ts->dg1, (size_t)b * 4096); want(&ts->dg2, (size_t)b * 4096); want(&ts->dg3, (size_t)b * 4096);
    want(&ts->R, (size_t)b * 36); want(&ts->dR, (size_t)b * 36); want(&ts->R64, (size_t)b * 16384); want(&ts->dR64, (size_t)b * 16384);
    for (int i = 0; i < 8; ++i) want(&ts->Ys[i], (size_t)b * 512 * 4);
    want(&ts->H1, (size_t)b * 2048); want(&ts->H2, (size_t)b * 1024);
    want(&ts->cls_logits, (size_t)b * C * 4); want(&ts->cls_prob, (size_t)b * C * 4);
    want(&ts->seg_logits, (size_t)M * P * 4); want(&ts->seg_prob, (size_t)M * P * 4);
    want(&ts->gvec, (size_t)b * 2048); want(&ts->dgvec, (size_t)b * 2048); want(&ts->fshift, 2048);
    want(&ts->colparts, colparts_doubles() * 8);
    want(&ts->G16, (size_t)M * 128 * 2);
    want(&ts->stage_x, (size_t)M * 12); want(&ts->stage_ycls, (size_t)b * 4); want(&ts->stage_yseg, (size_t)M * 4); want(&ts->stage_se3, (size_t)b * 36);
    want(&ts->step_scalars, 16);
    PCP_TRY(pcp_normalize_workspace_bytes(b, n, &ts->norm_ws_bytes));
    want(&ts->norm_ws, ts->norm_ws_bytes + 256);
    if (ts->graph_exec) { cudaGraphExecDestroy(ts->graph_exec); ts->graph_exec = nullptr; }
    ts->graph_seen = 0;
    for (int i = 0; i < 4; ++i) want(&ts->dsmall[i], (size_t)b * 4096 * 4);
    want(&ts->Y64a, M * 256); want(&ts->Y64b, M * 256); want(&ts->Y128, M * 512); want(&ts->Y128b, M * 512);
    want(&ts->Y256, M * 1024); want(&ts->Y512, M * 2048);
    if (!ts->implicit_pool) want(&ts->G1024, M * 4096);
    else
        for (PoolBufs &pb : ts->pool) {
            want(&pb.ext, (size_t)b * 1024 * 16); want(&pb.gram, 128 * 128 * 4); want(&pb.asum, 128 * 8); want(&pb.zsel, (size_t)b * 4096);
            want(&pb.gram_parts, tc_gram_scratch_floats(128) * 4); want(&pb.colparts, (size_t)sm_count() * 4 * 128 * 8);
            want(&pb.gram_g, 128 * 128 * 4); want(&pb.asum_g, 128 * 8);
            want(&pb.arg, (size_t)b * 4096); want(&pb.S, (size_t)b * 4096); want(&pb.c0, 4096); want(&pb.c1, 4096); want(&pb.Wt, 1024 * 128 * 4);
            want(&pb.Q, 128 * 128 * 4); want(&pb.q, 512);
        }
    want(&ts->G512, M * 2048); want(&ts->G256, M * 1024); want(&ts->G128a, M * 512); want(&ts->G128b, M * 512);
    want(&ts->G64a, M * 256); want(&ts->G64b, M * 256); want(&ts->GP, (size_t)M * P * 4);
    want(&ts->dX64, M * 256); want(&ts->dX2, M * 256); want(&ts->dX3, M * 12);
    want(&ts->metrics, 8 * 8);
    size_t total = 0;
    for (auto &r : req) total += r.second;
    PCP_CUDA_OK(cudaMalloc(&ts->arena, total));
    size_t off = 0;
    for (auto &r : req) { *r.first = ts->arena + off; off += r.second; }
    ts->arena_bytes = total; ts->b = b; ts->n = n;
    return PCP_OK;
}

// blocks per SM of the column reductions whose blocks end in one atomic per column (colstats4, bn_bwd_reduce4, wgrad3)
static inline int colred_waves()
{
    static const int w = [] { const char *e = getenv("PCP_COLRED_WAVES"); const int v = e ? atoi(e) : 0; return v >= 1 && v <= 8 ? v : 4; }();
    return w;
}

static inline int grid_for(int64_t work, int per_block = 256, int waves = 8)
{
    const int64_t g = (work + per_block - 1) / per_block;
    return (int)std::max<int64_t>(1, std::min<int64_t>(g, (int64_t)sm_count() * waves));
}

static int softmax_ce(const float *logits, int64_t rows, int C, const int32_t *labels, float gscale, float *prob, float *dlogits, double *loss_sum,
                      double *correct, cudaStream_t st)
{
    if (C <= 96) {
        const int grid = (int)std::max<int64_t>(1, std::min<int64_t>((rows + 127) / 128, (int64_t)sm_count() * 4));
        softmax_ce_tile_kernel<<<grid, 128, (size_t)128 * C * sizeof(float), st>>>(logits, rows, C, labels, gscale, prob, dlogits, loss_sum, correct);
    } else {
        softmax_ce_kernel<<<grid_for(rows), 256, 0, st>>>(logits, rows, C, labels, gscale, prob, dlogits, loss_sum, correct);
    }
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// ---- GEMM wrappers (row-major fp32 activations; W is (cin x cout) row-major inside the flat parameter buffer) ----
// PCP_PRECISION_BF16: operands are rounded to bf16 and multiplied on tcgen05 with fp32 accumulation (tc_gemm / tc_wgrad);
// shapes the tensor-core tiles do not cover (K = 3 inputs, 9 / C / P wide outputs as a contraction) stay on the fp32 path.
struct TcCtx {
    bool on = false;
    uint8_t *img = nullptr;       // scratch weight image (frozen layers with a folded BatchNorm, anything the cache cannot hold)
    size_t img_bytes = 0;
    ImgCache *cache = nullptr;
};
static thread_local TcCtx g_tc;
constexpr size_t kImgCacheBytes = (size_t)48 << 20;
constexpr int kImgCacheMaxJobs = 256;

// allocation and the (pageable-source) table upload: never inside a stream capture
static int img_cache_sync_table(ImgCache *c, cudaStream_t st)
{
    if (!c->pool) {
        PCP_CUDA_OK(cudaMalloc(&c->pool, kImgCacheBytes));
        PCP_CUDA_OK(cudaMalloc(&c->d_jobs, kImgCacheMaxJobs * sizeof(PackJob)));
        c->cap = kImgCacheBytes;
    }
    const int n = (int)c->jobs.size();
    if (n > c->d_count) {
        // the table only grows during the first steps; the copy is ordered before the launch on the same stream and
        // completes (pageable source) before the call returns
        PCP_CUDA_OK(cudaMemcpyAsync(c->d_jobs + c->d_count, c->jobs.data() + c->d_count, (size_t)(n - c->d_count) * sizeof(PackJob),
                                    cudaMemcpyHostToDevice, st));
        c->d_count = n;
    }
    return PCP_OK;
}
// re-pack every image of the device table (stream-ordered, capturable)
static int img_cache_refresh(ImgCache *c, cudaStream_t st)
{
    PCP_TRY(pack_weight_images(c->d_jobs, c->d_count, st));
    c->fresh = c->d_count;
    return PCP_OK;
}

// image of the logical (K x N) matrix B[k][n] = W[k*ldw + n] (or W[n*ldw + k], transpose) for tc_gemm
static int tc_image(const float *W, int K, int N, int ldw, bool transpose, const uint8_t **img, cudaStream_t st, const float *sgn = nullptr)
{
    ImgCache *c = g_tc.cache;
    const size_t bytes = (gemm_b_image_bytes(K, N) + 1023) & ~(size_t)1023;
    if (c && c->pool && W >= c->lo && W < c->hi && (!sgn || (sgn >= c->lo && sgn < c->hi))) {
        for (int i = 0; i < (int)c->jobs.size(); ++i) {
            const PackJob &j = c->jobs[i];
            if (j.W == W && j.K == K && j.N == N && j.ldw == ldw && j.transpose == (transpose ? 1 : 0) && j.sgn == sgn) {
                if (i >= c->fresh) PCP_TRY(pack_weight_image(W, K, N, ldw, transpose, reinterpret_cast<uint8_t *>(j.img), st, nullptr, sgn));
                *img = reinterpret_cast<const uint8_t *>(j.img);
                return PCP_OK;
            }
        }
        if ((int)c->jobs.size() < kImgCacheMaxJobs && c->used + bytes <= c->cap) {
            PackJob j;
            j.W = W; j.img = reinterpret_cast<uint16_t *>(c->pool + c->used); j.K = K; j.N = N; j.ldw = ldw; j.transpose = transpose ? 1 : 0;
            j.bnt = tc_gemm_bn_tile(N); j.sgn = sgn;
            c->used += bytes;
            c->jobs.push_back(j);
            PCP_TRY(pack_weight_image(W, K, N, ldw, transpose, reinterpret_cast<uint8_t *>(j.img), st, nullptr, sgn));
            *img = reinterpret_cast<const uint8_t *>(j.img);
            return PCP_OK;
        }
    }
    PCP_REQUIRE(bytes <= g_tc.img_bytes, PCP_ERR_UNSUPPORTED, "tc_image: %d x %d does not fit the scratch image", K, N);
    PCP_TRY(pack_weight_image(W, K, N, ldw, transpose, g_tc.img, st, nullptr, sgn));
    *img = g_tc.img;
    return PCP_OK;
}
// scratch of the column reductions' per-block partial sums (TrainState::colparts; colparts_doubles() doubles)
static thread_local double *g_colparts = nullptr;

// Data-parallel BatchNorm: every batch-statistics reduction of the step is summed over the ranks through the
// host's callback before it is used, so a data-parallel step computes the single-device step on the concatenated batch.
struct SyncCtx {
    pcp_allreduce_fn fn = nullptr;
    void *user = nullptr;
    double scale = 1.0;            // global batch / local batch: rows of the concatenated batch = local rows * scale
};
static thread_local SyncCtx g_sync;
static inline int sync_sum(void *d_buf, size_t count, int dtype, cudaStream_t st)
{
    if (!g_sync.fn) return PCP_OK;
    const int rc = g_sync.fn(g_sync.user, d_buf, count, dtype, reinterpret_cast<pcp_stream>(st));
    PCP_REQUIRE(rc == 0, PCP_ERR_INVALID, "BatchNorm all-reduce callback failed (%d)", rc);
    return PCP_OK;
}

// C (rows x cout) = A (rows x cin) . W  [+ bias]          groups > 1: per-group bias (stride cout), rows per group
static int mm_nn(const float *A, int64_t rows, int cin, const float *W, int ldw, int cout, float *Cc, const float *bias, cudaStream_t st,
                 int groups = 1, int64_t bias_stride = 0, const __nv_bfloat16 *A16 = nullptr)
{
    if (g_tc.on && cin % 64 == 0 && (cout <= 128 || cout % 128 == 0) && gemm_b_image_bytes(cin, cout) <= g_tc.img_bytes) {
        TcGemmParams p;
        PCP_TRY(tc_image(W, cin, cout, ldw, false, &p.Bimg, st));
        const int64_t rpg = rows / groups;
        p.A = A; p.a_f32 = 1; p.lda = cin; p.a_group_stride = rpg * cin;
        if (A16) { p.A = A16; p.a_f32 = 0; } p.bias = bias; p.bias_group_stride = bias_stride;
        p.C32 = Cc; p.ldc = cout; p.c_group_stride = rpg * cout; p.M = (int)rpg; p.N = cout; p.K = cin; p.groups = groups;
        return tc_gemm(p, st);
    }
    if (cin == 3 && groups == 1 && cout % 4 == 0 && ldw % 4 == 0 &&
        (((uintptr_t)W | (uintptr_t)Cc | (uintptr_t)(bias ? bias : W)) & 15u) == 0) {
        mm3_kernel<<<grid_for(rows * (cout / 4)), 256, 0, st>>>(A, W, ldw, bias, rows, cout / 4, reinterpret_cast<float4 *>(Cc));
        PCP_CHECK_LAUNCH();
        return PCP_OK;
    }
    GemmArgs g;
    g.A = A; g.lda = cin; g.B = W; g.ldb = ldw; g.C = Cc; g.ldc = cout; g.bias = bias; g.N = cout; g.K = cin;
    if (groups > 1) {
        const int64_t rpg = rows / groups;
        g.groups = groups; g.M = (int)rpg; g.strideA = rpg * cin; g.strideC = rpg * cout; g.strideBias = bias_stride;
    } else {
        g.M = (int)rows;
    }
    return gemm_f32(g, st);
}
// dA (rows x cin) (+)= dZ (rows x cout) . W^T
static int mm_nt(const float *dZ, int64_t rows, int cout, const float *W, int ldw, int cin, float *dA, bool accumulate, cudaStream_t st,
                 const __nv_bfloat16 *dZ16 = nullptr)
{
    if (g_tc.on && cout % 64 == 0 && (cin <= 128 || cin % 128 == 0) && gemm_b_image_bytes(cout, cin) <= g_tc.img_bytes) {
        TcGemmParams p;
        PCP_TRY(tc_image(W, cout, cin, ldw, true, &p.Bimg, st));
        p.A = dZ; p.a_f32 = 1; p.lda = cout; p.C32 = dA; p.ldc = cin; p.M = (int)rows; p.N = cin; p.K = cout;
        if (dZ16) { p.A = dZ16; p.a_f32 = 0; }      // the producer's own bf16 copy (bn_bwd): same rounding, persistent TMA kernel
        p.accumulate = accumulate ? 1 : 0;
        // few row tiles and a long contraction (the T-Net tails' 4096 -> 256): deal the k-tiles over CTAs
        const int kt = cout / 64;
        if (rows <= 512 && kt >= 16) {
            int ks = 16;
            while (kt % ks) ks >>= 1;
            if (ks > 1) {
                if (!accumulate) PCP_CUDA_OK(cudaMemsetAsync(dA, 0, (size_t)rows * cin * sizeof(float), st));
                p.ksplit = ks; p.accumulate = 0;
            }
        }
        return tc_gemm(p, st);
    }
    GemmArgs g;
    g.A = dZ; g.lda = cout; g.B = W; g.ldb = ldw; g.transB = 1; g.C = dA; g.ldc = cin; g.M = (int)rows; g.N = cin; g.K = cout;
    g.accumulate = accumulate ? 1 : 0;
    return gemm_f32(g, st);
}
// dW (cin x cout) += A^T (cin x rows) . dZ (rows x cout)   — split-K over the rows, atomics into the zeroed gradient buffer
static int mm_tn(const float *A, int64_t rows, int cin, const float *dZ, int cout, float *dW, int lddw, cudaStream_t st)
{
    if (g_tc.on && cin % 8 == 0 && cout % 8 == 0 && ((((uintptr_t)A | (uintptr_t)dZ) & 15u) == 0))
        return tc_wgrad(A, cin, rows, cin, dZ, cout, cout, dW, lddw, st);
    if (cin == 3 && lddw == cout && colmap4_ok(cout) && (((uintptr_t)dZ & 15u) == 0) && g_colparts) {
        const int np = grid_for(rows, 64, colred_waves());
        wgrad3_kernel<<<np, 256, 0, st>>>(A, reinterpret_cast<const float4 *>(dZ), rows, cout, g_colparts);
        PCP_CHECK_LAUNCH();
        reduce_parts_kernel<<<3 * cout, 256, 0, st>>>(g_colparts, np, 3 * cout, nullptr, nullptr, nullptr, 0, dW);
        PCP_CHECK_LAUNCH();
        return PCP_OK;
    }
    GemmArgs g;
    g.A = A; g.lda = cin; g.transA = 1; g.B = dZ; g.ldb = cout; g.C = dW; g.ldc = lddw; g.M = cin; g.N = cout; g.K = (int)rows;
    const int64_t blocks = (int64_t)((cin + 127) / 128) * ((cout + 63) / 64);
    int64_t sk = (4 * (int64_t)sm_count() + blocks - 1) / blocks;
    sk = std::min<int64_t>(sk, std::max<int64_t>(1, rows / 128));
    sk = std::max<int64_t>(1, std::min<int64_t>(sk, 4096));
    g.splitk = (int)sk;
    if (sk == 1) g.accumulate = 1;
    return gemm_f32(g, st);
}

#include "train_pool.inl"

static inline float *param(pcp_model *m, int ti) { return m->d_params + m->tensors[ti].offset; }
static inline float *grad(pcp_model *m, int ti) { return m->d_grads + m->tensors[ti].offset; }

// BN statistics + (mean, inv, sc, beta) for layer t whose Z is filled
static int bn_prepare(pcp_model *m, TL &t, cudaStream_t st)
{
    const LayerSpec &S = m->layers[t.id];
    const double *parts = t.red;
    int np = 1;
    if (t.batch) {
        np = grid_for(t.rows, 64, colred_waves());
        if (colmap4_ok(t.cout))
            colstats4_kernel<<<np, 256, 0, st>>>(reinterpret_cast<const float4 *>(t.Z), t.rows, t.cout, g_colparts);
        else
            colstats_kernel<<<np, 256, 0, st>>>(t.Z, t.rows, t.cout, g_colparts);
        PCP_CHECK_LAUNCH();
        parts = g_colparts;
        if (g_sync.fn) {             // the ranks exchange the reduced sums
            reduce_parts_kernel<<<2 * t.cout, 256, 0, st>>>(g_colparts, np, 2 * t.cout, t.red, nullptr, nullptr, 0, nullptr);
            PCP_CHECK_LAUNCH();
            PCP_TRY(sync_sum(t.red, (size_t)t.cout * 2, 1, st));
            parts = t.red; np = 1;
        }
    }
    bn_finalize_kernel<<<t.cout, 256, 0, st>>>(parts, np, (double)t.rows * g_sync.scale, t.cout, t.batch ? 1 : 0, param(m, S.t_gamma),
                                                        param(m, S.t_beta), param(m, S.t_mean), param(m, S.t_var), t.mean, t.inv, t.sc, t.beta,
                                                        kBnEpsT, kBnMomentum);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

static int bn_act(const TL &t, float *Y, cudaStream_t st, __nv_bfloat16 *Y16 = nullptr)
{
    const int64_t n4 = t.rows * t.cout / 4;
    bn_act_kernel<<<grid_for(n4), 256, 0, st>>>(reinterpret_cast<const float4 *>(t.Z), n4, t.cout / 4, t.bnp(), reinterpret_cast<float4 *>(Y),
                                                reinterpret_cast<uint2 *>(Y16));
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// the layer's activation relu(bn(Z)) in the backward pass: the forward pass's own copy, else recomputed into scratch
static int layer_act(const TL &t, float *scratch, const float **Y, cudaStream_t st)
{
    if (t.Y) { *Y = t.Y; return PCP_OK; }
    PCP_TRY(bn_act(t, scratch, st));
    *Y = scratch;
    return PCP_OK;
}

// forward of one BN layer: Z = A . W ; statistics ; optionally Y = relu(bn(Z))
// A16 (optional): A already rounded to bf16 by its producer; Y16 (optional): receives bf16(Y) for the next layer
static int layer_fwd_train(pcp_model *m, TL &t, const float *A, float *Y, cudaStream_t st, const __nv_bfloat16 *A16 = nullptr,
                           __nv_bfloat16 *Y16 = nullptr)
{
    const LayerSpec &S = m->layers[t.id];
    PCP_TRY(mm_nn(A, t.rows, t.cin, param(m, S.t_kernel), t.cout, t.cout, t.Z, nullptr, st, 1, 0, A16));
    PCP_TRY(bn_prepare(m, t, st));
    if (Y) PCP_TRY(bn_act(t, Y, st, Y16));
    return PCP_OK;
}

// X = pc @ R per cloud (PointNet.py:207), 3 x 3: one thread per point, the fma order of the generic GEMM it replaces
__global__ void __launch_bounds__(256)
pc_matmul3_kernel(const float *__restrict__ X, const float *__restrict__ R, int64_t n, int64_t total, float *__restrict__ Y)
{
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const float *r = R + (p / n) * 9;
        const float x0 = __ldg(X + 3 * p), x1 = __ldg(X + 3 * p + 1), x2 = __ldg(X + 3 * p + 2);
#pragma unroll
        for (int j = 0; j < 3; ++j) Y[3 * p + j] = fmaf(x2, __ldg(r + 6 + j), fmaf(x1, __ldg(r + 3 + j), x0 * __ldg(r + j)));
    }
}

// dR[g] (3 x 3) += X[g]^T (3 x n) . dX[g] (n x 3): one block per cloud (a 128 x 64 GEMM tile per cloud took 146 us for this)
__global__ void __launch_bounds__(256)
atb3_kernel(const float *__restrict__ X, const float *__restrict__ dX, int64_t n, float *__restrict__ dR)
{
    __shared__ float sh[8][9];
    const float *x = X + (int64_t)blockIdx.x * n * 3, *d = dX + (int64_t)blockIdx.x * n * 3;
    float acc[9] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int64_t r = threadIdx.x; r < n; r += 256) {
        const float xv[3] = {__ldg(x + 3 * r), __ldg(x + 3 * r + 1), __ldg(x + 3 * r + 2)};
        const float dv[3] = {__ldg(d + 3 * r), __ldg(d + 3 * r + 1), __ldg(d + 3 * r + 2)};
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) acc[i * 3 + j] = fmaf(xv[i], dv[j], acc[i * 3 + j]);
    }
#pragma unroll
    for (int e = 0; e < 9; ++e) {
#pragma unroll
        for (int o = 16; o; o >>= 1) acc[e] += __shfl_xor_sync(0xffffffffu, acc[e], o);
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5][e] = acc[e];
    }
    __syncthreads();
    if (threadIdx.x < 9) {
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) v += sh[w][threadIdx.x];
        dR[(int64_t)blockIdx.x * 9 + threadIdx.x] += v;
    }
}

// dR[g] (64 x 64) += X[g]^T (64 x n) . D[g] (n x 64) — the feature transform's gradient (X_64 = X @ R_64, PointNet.py:228).
// A CTA takes kRows rows of one cloud through shared memory in 64-row slabs; thread (ty, tx) owns the 4 x 4 block
// (X columns 4 ty .., D columns 4 tx ..): two 16-byte shared loads per 16 multiply-adds, the X one a broadcast.  The CTAs of a
// cloud add their blocks with atomics (the generic 128 x 64 GEMM tile wastes half its rows on a 64 x 64 result: 133 us).
constexpr int kAtbRows = 512;
__global__ void __launch_bounds__(256)
atb64_kernel(const float *__restrict__ X, const float *__restrict__ D, int64_t n, float *__restrict__ dR)
{
    __shared__ __align__(16) float xs[64][64], ds[64][64];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int64_t g = blockIdx.y, r0 = (int64_t)blockIdx.x * kAtbRows, r1 = min(n, r0 + kAtbRows);
    const float4 *X4 = reinterpret_cast<const float4 *>(X + (g * n) * 64), *D4 = reinterpret_cast<const float4 *>(D + (g * n) * 64);
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int64_t s0 = r0; s0 < r1; s0 += 64) {
        const int rows = (int)min((int64_t)64, r1 - s0);
        __syncthreads();
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int idx = tid + 256 * u, r = idx >> 4, c4 = idx & 15;       // 64 rows x 16 float4
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
            const float4 xv = r < rows ? __ldg(X4 + (s0 + r) * 16 + c4) : zero, dv = r < rows ? __ldg(D4 + (s0 + r) * 16 + c4) : zero;
            *reinterpret_cast<float4 *>(&xs[r][c4 * 4]) = xv;
            *reinterpret_cast<float4 *>(&ds[r][c4 * 4]) = dv;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < 64; ++k) {
            const float4 a = *reinterpret_cast<const float4 *>(&xs[k][ty * 4]), bq = *reinterpret_cast<const float4 *>(&ds[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {bq.x, bq.y, bq.z, bq.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
    }
    float *out = dR + g * 4096;
#pragma unroll
    for (int i = 0; i < 4; ++i)          // one 16-byte reduction per row of the thread's block (red.global.add.v4.f32)
        atomicAdd(reinterpret_cast<float4 *>(out + (ty * 4 + i) * 64 + tx * 4), make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
}

// out[g, c] = beta[c] - sc[c] * mean[c] (+ sc[c] * add[g, c]): what a frozen BatchNorm leaves of itself once sc is folded into
// the layer's weight image — relu(sc * (a . W + add - mean) + beta) = relu(a . (W sc) + out)
__global__ void __launch_bounds__(256)
frozen_shift_kernel(BnP p, int C, const float *__restrict__ add, int groups, float *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= groups * C) return;
    const int c = i % C;
    float v = fmaf(-p.sc[c], p.mean[c], p.beta[c]);
    if (add) v = fmaf(p.sc[c], add[i], v);
    out[i] = v;
}

// Forward of a BN layer that runs on its MOVING statistics (keras: not trainable) when nothing will be differentiated
// through it: BatchNorm and ReLU ride in the GEMM (scaled weight image, shift + relu epilogue), the output is bf16 — what
// the next GEMM reads anyway — and the pre-activation Z is never written.  A: rows x cin (fp32 or bf16), W: first cin rows
// of the layer's kernel; add: optional per-group pre-BN bias (groups x cout), shift: groups x cout floats of scratch.
static int frozen_layer_fwd16(pcp_model *m, TL &t, const void *A, int a_f32, int cin, const float *add, int groups, float *shift,
                              __nv_bfloat16 *Y, cudaStream_t st)
{
    const LayerSpec &S = m->layers[t.id];
    PCP_TRY(bn_prepare(m, t, st));
    frozen_shift_kernel<<<(groups * t.cout + 255) / 256, 256, 0, st>>>(t.bnp(), t.cout, add, groups, shift);
    PCP_CHECK_LAUNCH();
    PCP_REQUIRE(gemm_b_image_bytes(cin, t.cout) <= g_tc.img_bytes, PCP_ERR_UNSUPPORTED, "frozen layer: weight image too large");
    PCP_TRY(pack_weight_image(param(m, S.t_kernel), cin, t.cout, t.cout, false, g_tc.img, st, t.sc));
    TcGemmParams p;
    const int64_t rpg = t.rows / groups;
    p.A = A; p.a_f32 = a_f32; p.lda = cin; p.a_group_stride = rpg * cin; p.Bimg = g_tc.img; p.bias = shift;
    p.bias_group_stride = groups > 1 ? t.cout : 0; p.C16 = Y; p.ldc = t.cout; p.c_group_stride = rpg * t.cout;
    p.M = (int)rpg; p.N = t.cout; p.K = cin; p.groups = groups; p.relu = 1;
    return tc_gemm(p, st);
}

// reduce_max over the points of each cloud of relu(bn(Z))
static int pool_fwd(const TL &t, int b, int64_t n, float *g, cudaStream_t st)
{
    PCP_CUDA_OK(cudaMemsetAsync(g, 0, (size_t)b * t.cout * 4, st));
    const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((n + 63) / 64, std::max<int64_t>(1, (int64_t)sm_count() * 8 / b)));
    bn_relu_colmax_kernel<<<dim3(gx, b), 256, 0, st>>>(t.Z, n, t.cout, t.bnp(), g);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// G (rows x cout) holds dY on entry, dZ on exit; BN parameter gradients are written when the layer trains
static int bn_bwd(pcp_model *m, TL &t, float *G, bool relu, cudaStream_t st, __nv_bfloat16 *G16 = nullptr)
{
    const LayerSpec &S = m->layers[t.id];
    const int64_t n4 = t.rows * t.cout / 4;
    if (t.batch) {
        const int np = grid_for(t.rows, 64, colred_waves());
        if (colmap4_ok(t.cout))
            bn_bwd_reduce4_kernel<<<np, 256, 0, st>>>(reinterpret_cast<const float4 *>(G), reinterpret_cast<const float4 *>(t.Z), t.rows, t.cout, t.bnp(),
                                                     relu ? 1 : 0, g_colparts);
        else
            bn_bwd_reduce_kernel<<<np, 256, 0, st>>>(G, t.Z, t.rows, t.cout, t.bnp(), relu ? 1 : 0, g_colparts);
        PCP_CHECK_LAUNCH();
        // s1 | s2 into t.red, and this rank's dbeta = s1, dgamma = s2 when the layer trains
        reduce_parts_kernel<<<2 * t.cout, 256, 0, st>>>(g_colparts, np, 2 * t.cout, t.red, t.trainable ? grad(m, S.t_beta) : nullptr,
                                                                 t.trainable ? grad(m, S.t_gamma) : nullptr, t.cout, nullptr);
        PCP_CHECK_LAUNCH();
    }
    if (t.batch) PCP_TRY(sync_sum(t.red, (size_t)t.cout * 2, 1, st));          // after the LOCAL sums became dgamma / dbeta
    const float inv_rows = (float)(1.0 / ((double)t.rows * g_sync.scale));
    if (colmap4_ok(t.cout))
        bn_bwd_apply4_kernel<<<grid_for(n4, 512), 256, 0, st>>>(reinterpret_cast<float4 *>(G), reinterpret_cast<const float4 *>(t.Z), t.rows, t.cout, t.bnp(),
                                                               relu ? 1 : 0, t.batch ? 1 : 0, t.red, t.red + t.cout, inv_rows,
                                                               reinterpret_cast<uint2 *>(G16));
    else {
        bn_bwd_apply_kernel<<<grid_for(n4), 256, 0, st>>>(reinterpret_cast<float4 *>(G), reinterpret_cast<const float4 *>(t.Z), n4, t.cout / 4, t.bnp(),
                                                         relu ? 1 : 0, t.batch ? 1 : 0, t.red, t.red + t.cout, inv_rows);
        if (G16) to_bf16_kernel<<<grid_for(n4), 256, 0, st>>>(reinterpret_cast<const float4 *>(G), n4, reinterpret_cast<uint2 *>(G16));
    }
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// gradient of the per-cloud max-pool into G (M x cout), then BN backward in place
static int pool_bwd(pcp_model *m, TL &t, int b, int64_t n, const float *g, const float *dg, int *cnt, float *G, cudaStream_t st)
{
    PCP_CUDA_OK(cudaMemsetAsync(cnt, 0, (size_t)b * t.cout * 4, st));
    const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((n + 63) / 64, std::max<int64_t>(1, (int64_t)sm_count() * 8 / b)));
    pool_count_kernel<<<dim3(gx, b), 256, 0, st>>>(t.Z, n, t.cout, t.bnp(), g, cnt);
    PCP_CHECK_LAUNCH();
    const int64_t total4 = t.rows * t.cout / 4;
    pool_bwd_kernel<<<grid_for(total4), 256, 0, st>>>(reinterpret_cast<const float4 *>(t.Z), n * t.cout / 4, total4, t.cout / 4, t.bnp(), g, dg, cnt,
                                                     reinterpret_cast<float4 *>(G));
    PCP_CHECK_LAUNCH();
    return bn_bwd(m, t, G, true, st);
}

// weight gradient + input gradient of a layer whose dZ is in G
static int layer_bwd_mm(pcp_model *m, TL &t, const float *Ain, const float *G, float *dA, bool accumulate, cudaStream_t st,
                        const __nv_bfloat16 *G16 = nullptr)
{
    const LayerSpec &S = m->layers[t.id];
    if (t.trainable) PCP_TRY(mm_tn(Ain, t.rows, t.cin, G, t.cout, grad(m, S.t_kernel), t.cout, st));
    if (dA) PCP_TRY(mm_nt(G, t.rows, t.cout, param(m, S.t_kernel), t.cout, t.cin, dA, accumulate, st, G16));
    return PCP_OK;
}

static int colsum(const float *X, int groups, int64_t rpg, int C, float *out, cudaStream_t st)
{
    const int gz = (C + 255) / 256;
    const int gx = (int)std::max<int64_t>(1, std::min<int64_t>((rpg + 31) / 32, std::max<int64_t>(1, (int64_t)sm_count() * 4 / ((int64_t)groups * gz))));
    group_colsum_kernel<<<dim3(gx, groups, gz), 256, 0, st>>>(X, rpg, C, out);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// the pooled layer li as the implicit path sees it (train_pool.inl)
static PoolLayer make_pool_layer(pcp_model *m, TrainState *ts, int li, int slot)
{
    const LayerSpec &S = m->layers[li];
    const TL &t = ts->L[li];
    PoolLayer P;
    P.W = param(m, S.t_kernel); P.gamma = param(m, S.t_gamma); P.beta_in = param(m, S.t_beta);
    P.mov_mean = param(m, S.t_mean); P.mov_var = param(m, S.t_var);
    P.mean = t.mean; P.inv = t.inv; P.sc = t.sc; P.beta = t.beta; P.red = t.red;
    P.b = ts->b; P.n = ts->n; P.cout = t.cout; P.batch = t.batch; P.trainable = t.trainable; P.pb = ts->pool[slot];
    return P;
}

// pooled layer forward: g = reduce_max(relu(bn(A . W)))
static int pooled_layer_fwd(pcp_model *m, TrainState *ts, int li, int slot, const float *A, float *g, cudaStream_t st)
{
    if (ts->implicit_pool) return pooled_forward(make_pool_layer(m, ts, li, slot), A, g, st);
    PCP_TRY(layer_fwd_train(m, ts->L[li], A, nullptr, st));
    return pool_fwd(ts->L[li], ts->b, ts->n, g, st);
}

// pooled layer backward: dg -> parameter gradients and dA (rows x 128); Aprev is the layer whose activation feeds li
static int pooled_layer_bwd(pcp_model *m, TrainState *ts, int li, int slot, int lprev, const float *g, const float *dg, float *dA, cudaStream_t st)
{
    const LayerSpec &S = m->layers[li];
    const float *Aprev = nullptr;
    if (ts->implicit_pool) {
        PCP_TRY(layer_act(ts->L[lprev], ts->Y128, &Aprev, st));
        PoolLayer PL = make_pool_layer(m, ts, li, slot);
        PL.A16 = (g_tc.on && Aprev == ts->L[lprev].Y) ? ts->L[lprev].Y16 : nullptr;       // written by this step's forward bn_act
        return pooled_backward(PL, Aprev, g, dg, grad(m, S.t_kernel), grad(m, S.t_gamma), grad(m, S.t_beta), dA, st);
    }
    PCP_TRY(pool_bwd(m, ts->L[li], ts->b, ts->n, g, dg, ts->cnt, ts->G1024, st));
    PCP_TRY(layer_act(ts->L[lprev], ts->Y128, &Aprev, st));
    return layer_bwd_mm(m, ts->L[li], Aprev, ts->G1024, dA, false, st);
}

// T-Net tail forward (PointNet.py:431-442): g (b x 1024) -> dense 512 -> 256 -> (256 x k^2) + b
static int tnet_tail_fwd(pcp_model *m, TrainState *ts, int first, int t_w, int t_b, int k, const float *g, float *Y1, float *Y2, float *out,
                         cudaStream_t st)
{
    PCP_TRY(layer_fwd_train(m, ts->L[first + 3], g, Y1, st));
    PCP_TRY(layer_fwd_train(m, ts->L[first + 4], Y1, Y2, st));
    return mm_nn(Y2, ts->b, 256, param(m, t_w), k * k, k * k, out, param(m, t_b), st);
}

// T-Net tail backward: dM (b x k^2) -> dg (b x 1024)
static int tnet_tail_bwd(pcp_model *m, TrainState *ts, int first, int t_w, int t_b, int k, bool trainable, const float *g, const float *Y1,
                         const float *Y2, const float *dM, float *dg, cudaStream_t st)
{
    const int b = ts->b;
    float *d256 = ts->dsmall[1], *d512 = ts->dsmall[0];
    if (trainable) {
        PCP_TRY(mm_tn(Y2, b, 256, dM, k * k, grad(m, t_w), k * k, st));
        PCP_TRY(colsum(dM, 1, b, k * k, grad(m, t_b), st));
    }
    PCP_TRY(mm_nt(dM, b, k * k, param(m, t_w), k * k, 256, d256, false, st));
    PCP_TRY(bn_bwd(m, ts->L[first + 4], d256, true, st));
    PCP_TRY(layer_bwd_mm(m, ts->L[first + 4], Y1, d256, d512, false, st));
    PCP_TRY(bn_bwd(m, ts->L[first + 3], d512, true, st));
    PCP_TRY(layer_bwd_mm(m, ts->L[first + 3], g, d512, dg, false, st));
    return PCP_OK;
}

// Everything of the step that is ordered on the stream: forward, losses, backward into the flat gradient buffer.  No
// allocation, no host synchronisation, no host-memory copies — the body is what train_step_impl captures into a graph.
// dyn (device, optional): {dropout seed, sample offset} read by the dropout kernels instead of cfg's values.
static int train_step_body(pcp_model *m, TrainState *ts, const float *d_x, int b, int64_t n, const int32_t *y_cls, const int32_t *y_seg,
                           const float *y_se3, const pcp_train_config *cfg, int global_batch, const float *mask1, const float *mask2,
                           const unsigned long long *dyn, cudaStream_t st)
{
    if (g_tc.cache) PCP_TRY(img_cache_refresh(g_tc.cache, st));
    const int64_t M = (int64_t)b * n;
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    const bool full = !m->desc.vanilla;
    const bool tr_sh = m->trainable[PCP_GROUP_SHARED], tr_it = m->trainable[PCP_GROUP_INPUT_TRANSFORM] && full;
    const bool tr_cls = m->trainable[PCP_GROUP_CLS_HEAD], tr_seg = m->trainable[PCP_GROUP_SEG_HEAD];
    const float rate = m->desc.dropout_rate;
    const bool seg_needed = !(cfg->skip_unweighted_heads && cfg->w_seg == 0.f && !tr_seg);
    const bool do_seg_bwd = seg_needed && cfg->w_seg != 0.f && y_seg;
    const bool do_cls_bwd = cfg->w_cls != 0.f && y_cls;
    PCP_CUDA_OK(cudaMemsetAsync(m->d_grads, 0, m->n_params * sizeof(float), st));
    PCP_CUDA_OK(cudaMemsetAsync(ts->metrics, 0, 64, st));
    TL *L = ts->L;

    // =========================================== forward ===========================================
    PCP_TRY(pcp_normalize_ws_f32(d_x, b, n, ts->pcn, nullptr, nullptr, ts->norm_ws, ts->norm_ws_bytes, st));   // PointNet.py:202
    const float *X3 = ts->pcn;
    // tensor-core mode: a 64-wide activation is also written in bf16 by its producer for the GEMM right behind it
    // (a16: ts->G16, the backward pass's scratch; a16b: ts->Y64b, needed across two layers)
    __nv_bfloat16 *a16 = g_tc.on ? ts->G16 : nullptr, *a16b = g_tc.on ? reinterpret_cast<__nv_bfloat16 *>(ts->Y64b) : nullptr;
    if (full) {
        PCP_TRY(layer_fwd_train(m, L[L_IT_C1], ts->pcn, L[L_IT_C1].Y, st, nullptr, a16));                                 // TNet.call :423
        PCP_TRY(layer_fwd_train(m, L[L_IT_C2], L[L_IT_C1].Y, L[L_IT_C2].Y, st, a16, a16 ? L[L_IT_C2].Y16 : nullptr));                                // :424
        PCP_TRY(pooled_layer_fwd(m, ts, L_IT_C3, 0, L[L_IT_C2].Y, ts->g1, st));                             // :425-429
        PCP_TRY(tnet_tail_fwd(m, ts, L_IT_C1, m->t_it_w, m->t_it_b, 3, ts->g1, ts->Ys[0], ts->Ys[1], ts->R, st));
        pc_matmul3_kernel<<<grid_for(M), 256, 0, st>>>(ts->pcn, ts->R, n, M, ts->X3);                    // X = pc @ R  :207
        PCP_CHECK_LAUNCH();
        X3 = ts->X3;
    } else {
        PCP_TRY(model_eye3(ts->R, b, st));
    }
    PCP_TRY(layer_fwd_train(m, L[L_S1_1], X3, L[L_S1_1].Y, st, nullptr, a16));                                           // :217
    PCP_TRY(layer_fwd_train(m, L[L_S1_2], L[L_S1_1].Y, ts->X2, st, a16, (full && a16) ? a16b : nullptr));                                       // :220
    const float *X64 = ts->X2;
    if (full) {
        PCP_TRY(layer_fwd_train(m, L[L_FT_C1], ts->X2, L[L_FT_C1].Y, st, a16b, a16));
        PCP_TRY(layer_fwd_train(m, L[L_FT_C2], L[L_FT_C1].Y, L[L_FT_C2].Y, st, a16, a16 ? L[L_FT_C2].Y16 : nullptr));
        PCP_TRY(pooled_layer_fwd(m, ts, L_FT_C3, 1, L[L_FT_C2].Y, ts->g2, st));
        PCP_TRY(tnet_tail_fwd(m, ts, L_FT_C1, m->t_ft_w, m->t_ft_b, 64, ts->g2, ts->Ys[2], ts->Ys[3], ts->R64, st));   // :227
        GemmArgs g;                                                                                     // X_64 = X @ R_64 :228
        g.A = ts->X2; g.lda = 64; g.strideA = n * 64; g.B = ts->R64; g.ldb = 64; g.strideB = 4096;
        g.C = ts->X64; g.ldc = 64; g.strideC = n * 64; g.M = (int)n; g.N = 64; g.K = 64; g.groups = b;
        PCP_TRY(gemm_f32(g, st));
        X64 = ts->X64;
    }
    PCP_TRY(layer_fwd_train(m, L[L_S2_1], X64, L[L_S2_1].Y, st, nullptr, a16));                                          // :236
    PCP_TRY(layer_fwd_train(m, L[L_S2_2], L[L_S2_1].Y, L[L_S2_2].Y, st, a16, a16 ? L[L_S2_2].Y16 : nullptr));                                     // :239
    PCP_TRY(pooled_layer_fwd(m, ts, L_S2_3, 2, L[L_S2_2].Y, ts->g3, st));                                  // :242-248

    // ---- classification head (:252-263) ----
    const float *Yc1 = ts->Ys[4], *Yc2 = ts->Ys[5];
    PCP_TRY(layer_fwd_train(m, L[L_CLS1], ts->g3, ts->Ys[4], st));
    const bool use_drop = rate > 0.f;
    const float *H1 = Yc1, *H2 = Yc2;
    if (use_drop) {
        dropout_kernel<<<grid_for((int64_t)b * 512), 256, 0, st>>>(Yc1, b, 512, cfg->use_dropout_masks ? mask1 : nullptr, cfg->dropout_seed,
                                                                   cfg->sample_offset, 1u, rate, ts->H1, dyn);
        PCP_CHECK_LAUNCH();
        H1 = ts->H1;
    }
    PCP_TRY(layer_fwd_train(m, L[L_CLS2], H1, ts->Ys[5], st));
    if (use_drop) {
        dropout_kernel<<<grid_for((int64_t)b * 256), 256, 0, st>>>(Yc2, b, 256, cfg->use_dropout_masks ? mask2 : nullptr, cfg->dropout_seed,
                                                                   cfg->sample_offset, 2u, rate, ts->H2, dyn);
        PCP_CHECK_LAUNCH();
        H2 = ts->H2;
    }
    {
        const LayerSpec &S = m->layers[L_CLS3];
        PCP_TRY(mm_nn(H2, b, 256, param(m, S.t_kernel), C, C, ts->cls_logits, param(m, S.t_bias), st));
    }
    float *dlog_cls = ts->dsmall[2];
    PCP_REQUIRE(C <= 4096 && P <= 4096, PCP_ERR_UNSUPPORTED, "training supports up to 4096 classes / parts");
    if (y_cls) {
        PCP_TRY(softmax_ce(ts->cls_logits, b, C, y_cls, cfg->w_cls / (float)global_batch, ts->cls_prob, dlog_cls,
                                                               ts->metrics + 0, ts->metrics + 4, st));
    }

    // ---- segmentation head (:268-290): concat([X_64, tile(global)]) . W1 = X_64 . W1[0:64] + global . W1[64:1088] ----
    ts->have_seg = seg_needed;
    // nothing is differentiated through a frozen head whose loss is not weighted: it runs in its inference form (moving
    // statistics folded into the GEMMs, bf16 activations, no Z) — the classification_pretrain profile of BASELINE config 3
    bool seg_frozen_fwd = seg_needed && !do_seg_bwd && g_tc.on && (P <= 128);
    for (int i = L_SEG1; i <= L_SEG4; ++i) seg_frozen_fwd = seg_frozen_fwd && L[i].has_bn && !L[i].batch;
    if (seg_frozen_fwd) {
        const LayerSpec &S1 = m->layers[L_SEG1], &S5 = m->layers[L_SEG5];
        __nv_bfloat16 *y512 = reinterpret_cast<__nv_bfloat16 *>(ts->Y512), *y256 = reinterpret_cast<__nv_bfloat16 *>(ts->Y256),
                      *y128a = reinterpret_cast<__nv_bfloat16 *>(ts->Y128), *y128b = reinterpret_cast<__nv_bfloat16 *>(ts->Y128b);
        PCP_TRY(mm_nn(ts->g3, b, 1024, param(m, S1.t_kernel) + (size_t)64 * 512, 512, 512, ts->gvec, nullptr, st));
        // X_64 as bf16 first: the 64 -> 512 layer then streams it through TMA instead of converting it once per column block
        __nv_bfloat16 *x16 = reinterpret_cast<__nv_bfloat16 *>(ts->Y64b);
        to_bf16_kernel<<<grid_for(M * 16), 256, 0, st>>>(reinterpret_cast<const float4 *>(X64), M * 16, reinterpret_cast<uint2 *>(x16));
        PCP_CHECK_LAUNCH();
        PCP_TRY(frozen_layer_fwd16(m, L[L_SEG1], x16, 0, 64, ts->gvec, b, ts->dgvec, y512, st));
        PCP_TRY(frozen_layer_fwd16(m, L[L_SEG2], y512, 0, 512, nullptr, 1, ts->fshift, y256, st));
        PCP_TRY(frozen_layer_fwd16(m, L[L_SEG3], y256, 0, 256, nullptr, 1, ts->fshift, y128a, st));
        PCP_TRY(frozen_layer_fwd16(m, L[L_SEG4], y128a, 0, 128, nullptr, 1, ts->fshift, y128b, st));
        TcGemmParams p;
        PCP_TRY(tc_image(param(m, S5.t_kernel), 128, P, P, false, &p.Bimg, st));
        p.A = y128b; p.a_f32 = 0; p.lda = 128; p.bias = param(m, S5.t_bias); p.C32 = ts->seg_logits; p.ldc = P;
        p.M = (int)M; p.N = P; p.K = 128;
        PCP_TRY(tc_gemm(p, st));
        if (y_seg) {
            PCP_TRY(softmax_ce(ts->seg_logits, M, P, y_seg, 0.f, ts->seg_prob, nullptr, ts->metrics + 1, ts->metrics + 5, st));
        }
    } else if (seg_needed) {
        const LayerSpec &S1 = m->layers[L_SEG1];
        PCP_TRY(mm_nn(ts->g3, b, 1024, param(m, S1.t_kernel) + (size_t)64 * 512, 512, 512, ts->gvec, nullptr, st));
        TL &t1 = L[L_SEG1];
        PCP_TRY(mm_nn(X64, M, 64, param(m, S1.t_kernel), 512, 512, t1.Z, ts->gvec, st, b, 512));
        PCP_TRY(bn_prepare(m, t1, st));
        PCP_TRY(bn_act(t1, ts->Y512, st));
        PCP_TRY(layer_fwd_train(m, L[L_SEG2], ts->Y512, ts->Y256, st));
        PCP_TRY(layer_fwd_train(m, L[L_SEG3], ts->Y256, ts->Y128, st));
        PCP_TRY(layer_fwd_train(m, L[L_SEG4], ts->Y128, ts->Y128b, st));
        const LayerSpec &S5 = m->layers[L_SEG5];
        PCP_TRY(mm_nn(ts->Y128b, M, 128, param(m, S5.t_kernel), P, P, ts->seg_logits, param(m, S5.t_bias), st));
        if (y_seg) {
            PCP_TRY(softmax_ce(ts->seg_logits, M, P, y_seg, cfg->w_seg / ((float)global_batch * (float)n), ts->seg_prob,
                                                           do_seg_bwd ? ts->GP : nullptr, ts->metrics + 1, ts->metrics + 5, st));
        }
    }

    // =========================================== backward ===========================================
    PCP_CUDA_OK(cudaMemsetAsync(ts->dg3, 0, (size_t)b * 4096, st));
    PCP_CUDA_OK(cudaMemsetAsync(ts->dX64, 0, (size_t)M * 256, st));
    PCP_CUDA_OK(cudaMemsetAsync(ts->dR, 0, (size_t)b * 36, st));
    PCP_CUDA_OK(cudaMemsetAsync(ts->dR64, 0, (size_t)b * 16384, st));
    const bool trunk_bwd = tr_sh || tr_it;

    if (do_cls_bwd && (tr_cls || trunk_bwd)) {
        const LayerSpec &S3 = m->layers[L_CLS3];
        float *d256 = ts->dsmall[1], *d512 = ts->dsmall[0];
        if (tr_cls) {
            PCP_TRY(mm_tn(H2, b, 256, dlog_cls, C, grad(m, S3.t_kernel), C, st));
            PCP_TRY(colsum(dlog_cls, 1, b, C, grad(m, S3.t_bias), st));
        }
        PCP_TRY(mm_nt(dlog_cls, b, C, param(m, S3.t_kernel), C, 256, d256, false, st));
        if (use_drop) {
            dropout_kernel<<<grid_for((int64_t)b * 256), 256, 0, st>>>(d256, b, 256, cfg->use_dropout_masks ? mask2 : nullptr, cfg->dropout_seed,
                                                                       cfg->sample_offset, 2u, rate, d256, dyn);
            PCP_CHECK_LAUNCH();
        }
        PCP_TRY(bn_bwd(m, L[L_CLS2], d256, true, st));
        PCP_TRY(layer_bwd_mm(m, L[L_CLS2], H1, d256, d512, false, st));
        if (use_drop) {
            dropout_kernel<<<grid_for((int64_t)b * 512), 256, 0, st>>>(d512, b, 512, cfg->use_dropout_masks ? mask1 : nullptr, cfg->dropout_seed,
                                                                       cfg->sample_offset, 1u, rate, d512, dyn);
            PCP_CHECK_LAUNCH();
        }
        PCP_TRY(bn_bwd(m, L[L_CLS1], d512, true, st));
        PCP_TRY(layer_bwd_mm(m, L[L_CLS1], ts->g3, d512, trunk_bwd ? ts->dg3 : nullptr, true, st));
    }

    if (do_seg_bwd && (tr_seg || trunk_bwd)) {
        const LayerSpec &S5 = m->layers[L_SEG5], &S1 = m->layers[L_SEG1];
        // Y128b (seg_l4 output) is still live from the forward pass; the others are recomputed from Z
        if (tr_seg) {
            PCP_TRY(mm_tn(ts->Y128b, M, 128, ts->GP, P, grad(m, S5.t_kernel), P, st));
            PCP_TRY(colsum(ts->GP, 1, M, P, grad(m, S5.t_bias), st));
        }
        PCP_TRY(mm_nt(ts->GP, M, P, param(m, S5.t_kernel), P, 128, ts->G128a, false, st));
        PCP_TRY(bn_bwd(m, L[L_SEG4], ts->G128a, true, st));
        PCP_TRY(bn_act(L[L_SEG3], ts->Y128, st));
        PCP_TRY(layer_bwd_mm(m, L[L_SEG4], ts->Y128, ts->G128a, ts->G128b, false, st));
        PCP_TRY(bn_bwd(m, L[L_SEG3], ts->G128b, true, st));
        PCP_TRY(bn_act(L[L_SEG2], ts->Y256, st));
        PCP_TRY(layer_bwd_mm(m, L[L_SEG3], ts->Y256, ts->G128b, ts->G256, false, st));
        PCP_TRY(bn_bwd(m, L[L_SEG2], ts->G256, true, st));
        PCP_TRY(bn_act(L[L_SEG1], ts->Y512, st));
        PCP_TRY(layer_bwd_mm(m, L[L_SEG2], ts->Y512, ts->G256, ts->G512, false, st));
        PCP_TRY(bn_bwd(m, L[L_SEG1], ts->G512, true, st));
        // per-cloud sums of dZ1: gradient of the hoisted global term
        PCP_CUDA_OK(cudaMemsetAsync(ts->dgvec, 0, (size_t)b * 2048, st));
        PCP_TRY(colsum(ts->G512, b, n, 512, ts->dgvec, st));
        if (tr_seg) {
            PCP_TRY(mm_tn(X64, M, 64, ts->G512, 512, grad(m, S1.t_kernel), 512, st));
            PCP_TRY(mm_tn(ts->g3, b, 1024, ts->dgvec, 512, grad(m, S1.t_kernel) + (size_t)64 * 512, 512, st));
        }
        if (trunk_bwd) {
            PCP_TRY(mm_nt(ts->G512, M, 512, param(m, S1.t_kernel), 512, 64, ts->dX64, true, st));
            PCP_TRY(mm_nt(ts->dgvec, b, 512, param(m, S1.t_kernel) + (size_t)64 * 512, 512, 1024, ts->dg3, true, st));
        }
    }

    // rotation loss + regularisers act on R / R64 directly
    if (y_se3) {
        mse_kernel<<<grid_for((int64_t)b * 9, 256, 1), 256, 0, st>>>(ts->R, y_se3, (int64_t)b * 9, cfg->w_rot / ((float)global_batch * 9.f),
                                                                    (tr_it && cfg->w_rot != 0.f) ? ts->dR : nullptr, ts->metrics + 2);
        PCP_CHECK_LAUNCH();
    }
    if (full && m->desc.regularize_input_transform) {
        tnet_reg_kernel<<<b, 32, 2 * 9 * sizeof(float), st>>>(ts->R, 3, tr_it ? ts->dR : nullptr, ts->metrics + 3);
        PCP_CHECK_LAUNCH();
    }
    if (full && m->desc.regularize_feature_transform) {
        tnet_reg_kernel<<<b, 256, 2 * 4096 * sizeof(float), st>>>(ts->R64, 64, tr_sh ? ts->dR64 : nullptr, ts->metrics + 3);
        PCP_CHECK_LAUNCH();
    }

    if (trunk_bwd) {
        // ---- mlp_2 (:234-248) ----
        PCP_TRY(pooled_layer_bwd(m, ts, L_S2_3, 2, L_S2_2, ts->g3, ts->dg3, ts->G128a, st));
        // g16: bn_bwd also leaves dZ in bf16 for the input-gradient GEMM behind it (tensor-core mode)
        __nv_bfloat16 *g16 = g_tc.on ? ts->G16 : nullptr;
        PCP_TRY(bn_bwd(m, L[L_S2_2], ts->G128a, true, st, g16));
        PCP_TRY(layer_bwd_mm(m, L[L_S2_2], L[L_S2_1].Y, ts->G128a, ts->G64a, false, st, g16));
        PCP_TRY(bn_bwd(m, L[L_S2_1], ts->G64a, true, st, g16));
        PCP_TRY(layer_bwd_mm(m, L[L_S2_1], X64, ts->G64a, ts->dX64, true, st, g16));
        float *dX2 = ts->dX64;                     // vanilla: X_64 is X (:231)
        if (full) {
            // X_64 = X @ R_64: dR_64 = X^T . dX_64 per cloud ; dX = dX_64 . R_64^T
            // atomics into the zero-initialised (+ regulariser) dR64, like every other weight-gradient reduction of the step
            PCP_REQUIRE(b <= 65535, PCP_ERR_UNSUPPORTED, "training: more than 65535 clouds per step");
            atb64_kernel<<<dim3((unsigned)((n + kAtbRows - 1) / kAtbRows), (unsigned)b), 256, 0, st>>>(ts->X2, ts->dX64, n, ts->dR64);
            PCP_CHECK_LAUNCH();
            GemmArgs h;
            h.A = ts->dX64; h.lda = 64; h.strideA = n * 64; h.B = ts->R64; h.ldb = 64; h.strideB = 4096; h.transB = 1;
            h.C = ts->dX2; h.ldc = 64; h.strideC = n * 64; h.M = (int)n; h.N = 64; h.K = 64; h.groups = b;
            PCP_TRY(gemm_f32(h, st));
            dX2 = ts->dX2;
            // ---- feature T-Net ----
            PCP_TRY(tnet_tail_bwd(m, ts, L_FT_C1, m->t_ft_w, m->t_ft_b, 64, tr_sh, ts->g2, ts->Ys[2], ts->Ys[3], ts->dR64, ts->dg2, st));
            PCP_TRY(pooled_layer_bwd(m, ts, L_FT_C3, 1, L_FT_C2, ts->g2, ts->dg2, ts->G128a, st));
            PCP_TRY(bn_bwd(m, L[L_FT_C2], ts->G128a, true, st, g16));
            PCP_TRY(layer_bwd_mm(m, L[L_FT_C2], L[L_FT_C1].Y, ts->G128a, ts->G64a, false, st, g16));
            PCP_TRY(bn_bwd(m, L[L_FT_C1], ts->G64a, true, st, g16));
            PCP_TRY(layer_bwd_mm(m, L[L_FT_C1], ts->X2, ts->G64a, dX2, true, st, g16));
        }
        // ---- mlp_1 (:215-223): dX2 holds dY of s1_l2 ----
        PCP_TRY(bn_bwd(m, L[L_S1_2], dX2, true, st, g16));
        PCP_TRY(layer_bwd_mm(m, L[L_S1_2], L[L_S1_1].Y, dX2, ts->G64b, false, st, g16));
        PCP_TRY(bn_bwd(m, L[L_S1_1], ts->G64b, true, st));
        PCP_TRY(layer_bwd_mm(m, L[L_S1_1], X3, ts->G64b, (full && tr_it) ? ts->dX3 : nullptr, false, st));
        if (full && tr_it) {
            // X = pc @ R: dR += pc^T . dX per cloud
            atb3_kernel<<<b, 256, 0, st>>>(ts->pcn, ts->dX3, n, ts->dR);
            PCP_CHECK_LAUNCH();
        }
    }
    if (full && tr_it) {
        // ---- input T-Net: gradient enters through R only ----
        PCP_TRY(tnet_tail_bwd(m, ts, L_IT_C1, m->t_it_w, m->t_it_b, 3, true, ts->g1, ts->Ys[0], ts->Ys[1], ts->dR, ts->dg1, st));
        PCP_TRY(pooled_layer_bwd(m, ts, L_IT_C3, 0, L_IT_C2, ts->g1, ts->dg1, ts->G128a, st));
        __nv_bfloat16 *g16 = g_tc.on ? ts->G16 : nullptr;
        PCP_TRY(bn_bwd(m, L[L_IT_C2], ts->G128a, true, st, g16));
        PCP_TRY(layer_bwd_mm(m, L[L_IT_C2], L[L_IT_C1].Y, ts->G128a, ts->G64a, false, st, g16));
        PCP_TRY(bn_bwd(m, L[L_IT_C1], ts->G64a, true, st));
        PCP_TRY(layer_bwd_mm(m, L[L_IT_C1], ts->pcn, ts->G64a, nullptr, false, st));
    }
    return PCP_OK;
}

static int train_step_impl(pcp_model *m, const float *d_x, int b, int64_t n, const int32_t *y_cls, const int32_t *y_seg, const float *y_se3,
                           const pcp_train_config *cfg, int global_batch, const float *mask1, const float *mask2, float *h_metrics,
                           cudaStream_t st)
{
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    PCP_TRY(ensure_arena(m, ts, b, n));
    const size_t kImgBytes = (size_t)4 << 20;
    if (!ts->tc_img) PCP_CUDA_OK(cudaMalloc(&ts->tc_img, kImgBytes));
    g_tc.on = m->desc.precision == PCP_PRECISION_BF16;
    g_tc.img = ts->tc_img; g_tc.img_bytes = kImgBytes;
    static const bool no_img_cache = getenv("PCP_TRAIN_NO_IMG_CACHE") != nullptr;
    g_tc.cache = (g_tc.on && !no_img_cache) ? &ts->img_cache : nullptr;
    if (g_tc.cache) {
        g_tc.cache->lo = m->d_params; g_tc.cache->hi = m->d_params + m->n_params;
        PCP_TRY(img_cache_sync_table(g_tc.cache, st));
    }
    g_colparts = ts->colparts;
    g_sync = SyncCtx();
    if (ts->sync_fn && global_batch > b) { g_sync.fn = ts->sync_fn; g_sync.user = ts->sync_user; g_sync.scale = (double)global_batch / (double)b; }
    const bool tr_seg = m->trainable[PCP_GROUP_SEG_HEAD];
    const bool seg_needed = !(cfg->skip_unweighted_heads && cfg->w_seg == 0.f && !tr_seg);
    for (int i = 0; i < L_COUNT; ++i) {
        TL &t = ts->L[i];
        if (t.id < 0) continue;
        t.trainable = m->trainable[m->layers[i].group];
        t.batch = t.has_bn && t.trainable;                       // keras: batch statistics only when training AND trainable
    }

    // ---- the step as one graph launch.  Eligible: no host callback inside the step (sync-BN), no caller-owned dropout masks;
    // the same configuration has run eagerly twice (every lazily created thing exists, the weight-image table is complete).
    // The graph reads its inputs from staging buffers and the per-step dropout scalars from device memory.
    // PCP_TRAIN_GRAPH=0 keeps every step eager.
    const char *genv = getenv("PCP_TRAIN_GRAPH");
    const bool graph_ok = !(genv && genv[0] == '0') && !ts->graph_broken && !g_sync.fn && !cfg->use_dropout_masks;
    TrainState::StepKey key;
    memset(&key, 0, sizeof(key));                                  // padding bytes too: the keys are compared with memcmp
    key.b = b; key.n = n; key.global_batch = global_batch; key.w_cls = cfg->w_cls; key.w_seg = cfg->w_seg; key.w_rot = cfg->w_rot;
    key.skip = cfg->skip_unweighted_heads;
    for (int g = 0; g < 4; ++g) key.tr[g] = m->trainable[g] ? 1 : 0;
    key.has_cls = y_cls != nullptr; key.has_seg = y_seg != nullptr; key.has_se3 = y_se3 != nullptr;
    key.jobs = g_tc.cache ? (int)g_tc.cache->jobs.size() : 0;
    if (memcmp(&key, &ts->graph_key, sizeof(key)) != 0) {
        if (ts->graph_exec) { cudaGraphExecDestroy(ts->graph_exec); ts->graph_exec = nullptr; }
        memcpy(&ts->graph_key, &key, sizeof(key)); ts->graph_seen = 0;
    }
    bool ran = false;
    if (graph_ok && ts->graph_seen >= 2) {
        const int64_t M = (int64_t)b * n;
        if (!ts->graph_stream) {
            PCP_CUDA_OK(cudaStreamCreateWithFlags(&ts->graph_stream, cudaStreamNonBlocking));
            PCP_CUDA_OK(cudaEventCreateWithFlags(&ts->graph_in, cudaEventDisableTiming));
            PCP_CUDA_OK(cudaEventCreateWithFlags(&ts->graph_out, cudaEventDisableTiming));
        }
        cudaStream_t gs = ts->graph_stream;
        PCP_CUDA_OK(cudaEventRecord(ts->graph_in, st));            // fork: everything queued on the caller's stream comes first
        PCP_CUDA_OK(cudaStreamWaitEvent(gs, ts->graph_in, 0));
        PCP_CUDA_OK(cudaMemcpyAsync(ts->stage_x, d_x, (size_t)M * 12, cudaMemcpyDeviceToDevice, gs));
        if (y_cls) PCP_CUDA_OK(cudaMemcpyAsync(ts->stage_ycls, y_cls, (size_t)b * 4, cudaMemcpyDeviceToDevice, gs));
        if (y_seg) PCP_CUDA_OK(cudaMemcpyAsync(ts->stage_yseg, y_seg, (size_t)M * 4, cudaMemcpyDeviceToDevice, gs));
        if (y_se3) PCP_CUDA_OK(cudaMemcpyAsync(ts->stage_se3, y_se3, (size_t)b * 36, cudaMemcpyDeviceToDevice, gs));
        set_step_scalars_kernel<<<1, 1, 0, gs>>>(ts->step_scalars, (unsigned long long)cfg->dropout_seed, (long long)cfg->sample_offset);
        PCP_CHECK_LAUNCH();
        if (!ts->graph_exec) {
            cudaGraph_t graph = nullptr;
            cudaError_t e = cudaStreamBeginCapture(gs, cudaStreamCaptureModeThreadLocal);
            int rc = PCP_ERR_CUDA;
            if (e == cudaSuccess) {
                rc = train_step_body(m, ts, ts->stage_x, b, n, y_cls ? ts->stage_ycls : nullptr, y_seg ? ts->stage_yseg : nullptr,
                                     y_se3 ? ts->stage_se3 : nullptr, cfg, global_batch, nullptr, nullptr, ts->step_scalars, gs);
                e = cudaStreamEndCapture(gs, &graph);
            }
            if (rc == PCP_OK && e == cudaSuccess && graph) e = cudaGraphInstantiate(&ts->graph_exec, graph, 0);
            if (graph) cudaGraphDestroy(graph);
            if (rc != PCP_OK || e != cudaSuccess || !ts->graph_exec) {
                if (getenv("PCP_TRAIN_GRAPH_DEBUG"))
                    fprintf(stderr, "pcp: training-step capture failed (rc %d, %s; %s): steps stay eager\n", rc, cudaGetErrorString(e), pcp_last_error());
                cudaGetLastError();                                  // the capture did not take: this model stays eager
                ts->graph_exec = nullptr; ts->graph_broken = true;
            }
        }
        if (ts->graph_exec) {
            PCP_CUDA_OK(cudaGraphLaunch(ts->graph_exec, gs));
            ran = true;
            if (getenv("PCP_TRAIN_GRAPH_DEBUG")) fprintf(stderr, "pcp: training step replayed from its graph\n");
        }
        PCP_CUDA_OK(cudaEventRecord(ts->graph_out, gs));           // join: the caller's stream continues behind the step
        PCP_CUDA_OK(cudaStreamWaitEvent(st, ts->graph_out, 0));
    }
    if (!ran) {
        PCP_TRY(train_step_body(m, ts, d_x, b, n, y_cls, y_seg, y_se3, cfg, global_batch, mask1, mask2, nullptr, st));
        ts->graph_seen++;
    }
    m->dev_newer = true;          // BN moving statistics were updated on the device
    m->prepared = false;

    if (h_metrics) {
        double hm[8];
        PCP_CUDA_OK(cudaMemcpyAsync(hm, ts->metrics, 64, cudaMemcpyDeviceToHost, st));
        PCP_CUDA_OK(cudaStreamSynchronize(st));
        const double cls_loss = y_cls ? hm[0] / b : 0.0, seg_loss = (seg_needed && y_seg) ? hm[1] / ((double)b * n) : 0.0;
        const double se3 = y_se3 ? hm[2] / ((double)b * 9) : 0.0, reg = hm[3];
        h_metrics[1] = (float)cls_loss; h_metrics[2] = (float)seg_loss; h_metrics[3] = (float)se3;
        h_metrics[4] = y_cls ? (float)(hm[4] / b) : 0.f;
        h_metrics[5] = (seg_needed && y_seg) ? (float)(hm[5] / ((double)b * n)) : 0.f;
        h_metrics[6] = (float)sqrt(se3);
        h_metrics[7] = (float)reg;
        h_metrics[0] = (float)(cfg->w_cls * cls_loss + cfg->w_seg * seg_loss + cfg->w_rot * se3 + reg);
    }
    return PCP_OK;
}

}  // namespace pcp

using namespace pcp;

// test hook (tests only): the BatchNorm all-reduce callback + row scale the next pcp_selftest_pooled_layer calls run under
// (emulates the data-parallel protocol of the pooled layer on one device); fn == null clears it
static SyncCtx s_selftest_sync;
extern "C" int pcp_selftest_set_bn_allreduce(pcp_allreduce_fn fn, void *user, double row_scale)
{
    s_selftest_sync = SyncCtx();
    if (fn) { s_selftest_sync.fn = fn; s_selftest_sync.user = user; s_selftest_sync.scale = row_scale; }
    return PCP_OK;
}

// test hook (tests only): one implicitly handled max-pooled layer, forward + backward, host in / host out.
// A (b*n x 128), W (128 x cout), gamma / beta (cout), dg (b x cout) -> g (b x cout), mean / inv (cout), dW (128 x cout),
// dgamma / dbeta (cout), dA (b*n x 128).  batch != 0: BN uses batch statistics (moving mean 0 / variance 1 otherwise).
extern "C" int pcp_selftest_pooled_layer(const float *h_a, const float *h_w, const float *h_gamma, const float *h_beta, const float *h_dg, int b,
                                         int64_t n, int cout, int batch, float *h_g, float *h_mean, float *h_inv, float *h_dw, float *h_dgamma,
                                         float *h_dbeta, float *h_da)
{
    PCP_REQUIRE(h_a && h_w && h_gamma && h_beta && h_dg && h_g && h_mean && h_inv && h_dw && h_dgamma && h_dbeta && h_da && b >= 1 && n >= 1 &&
                    cout >= 128 && cout % 128 == 0 && cout <= 1024,
                PCP_ERR_INVALID, "pcp_selftest_pooled_layer: bad arguments");
    const int64_t rows = (int64_t)b * n;
    const size_t bc = (size_t)b * cout;
    std::vector<std::pair<void **, size_t>> req;
    auto want = [&](auto **p, size_t bytes) { req.emplace_back(reinterpret_cast<void **>(p), (bytes + 255) & ~(size_t)255); };
    float *dA_in, *dW_in, *gam, *bet, *mm, *mv, *dg, *g, *dWg, *dgam, *dbet, *dAo;
    uint8_t *img;
    PoolLayer L;
    want(&dA_in, rows * 512); want(&dW_in, (size_t)128 * cout * 4); want(&gam, cout * 4); want(&bet, cout * 4); want(&mm, cout * 4); want(&mv, cout * 4);
    want(&dg, bc * 4); want(&g, bc * 4); want(&dWg, (size_t)128 * cout * 4); want(&dgam, cout * 4); want(&dbet, cout * 4); want(&dAo, rows * 512);
    want(&img, (size_t)4 << 20);
    want(&L.mean, cout * 4); want(&L.inv, cout * 4); want(&L.sc, cout * 4); want(&L.beta, cout * 4); want(&L.red, (size_t)cout * 16);
    want(&L.pb.gram_parts, tc_gram_scratch_floats(128) * 4); want(&L.pb.colparts, (size_t)sm_count() * 4 * 128 * 8);
    want(&L.pb.ext, bc * 16); want(&L.pb.gram, 128 * 128 * 4); want(&L.pb.asum, 128 * 8); want(&L.pb.gram_g, 128 * 128 * 4); want(&L.pb.asum_g, 128 * 8); want(&L.pb.zsel, bc * 4); want(&L.pb.arg, bc * 4);
    want(&L.pb.S, bc * 4); want(&L.pb.c0, cout * 4); want(&L.pb.c1, cout * 4); want(&L.pb.Wt, (size_t)cout * 512); want(&L.pb.Q, 128 * 128 * 4);
    want(&L.pb.q, 512);
    size_t total = 0;
    for (auto &r : req) total += r.second;
    char *arena = nullptr;
    PCP_CUDA_OK(cudaMalloc(&arena, total));
    size_t off = 0;
    for (auto &r : req) { *r.first = arena + off; off += r.second; }
    int rc = PCP_OK;
    auto run = [&]() -> int {
        PCP_CUDA_OK(cudaMemset(arena, 0, total));
        PCP_CUDA_OK(cudaMemcpy(dA_in, h_a, rows * 512, cudaMemcpyHostToDevice));
        PCP_CUDA_OK(cudaMemcpy(dW_in, h_w, (size_t)128 * cout * 4, cudaMemcpyHostToDevice));
        PCP_CUDA_OK(cudaMemcpy(gam, h_gamma, cout * 4, cudaMemcpyHostToDevice));
        PCP_CUDA_OK(cudaMemcpy(bet, h_beta, cout * 4, cudaMemcpyHostToDevice));
        PCP_CUDA_OK(cudaMemcpy(dg, h_dg, bc * 4, cudaMemcpyHostToDevice));
        std::vector<float> ones(cout, 1.f);
        PCP_CUDA_OK(cudaMemcpy(mv, ones.data(), cout * 4, cudaMemcpyHostToDevice));
        g_tc = TcCtx();                    // thread-local: a training step on this thread leaves its own context behind
        g_tc.on = true; g_tc.img = img; g_tc.img_bytes = (size_t)4 << 20;
        g_sync = SyncCtx();
        if (s_selftest_sync.fn) g_sync = s_selftest_sync;
        L.W = dW_in; L.gamma = gam; L.beta_in = bet; L.mov_mean = mm; L.mov_var = mv; L.b = b; L.n = n; L.cout = cout;
        L.batch = batch != 0; L.trainable = true;
        PCP_TRY(pooled_forward(L, dA_in, g, nullptr));
        PCP_TRY(pooled_backward(L, dA_in, g, dg, dWg, dgam, dbet, dAo, nullptr));
        PCP_CUDA_OK(cudaDeviceSynchronize());
        PCP_CUDA_OK(cudaMemcpy(h_g, g, bc * 4, cudaMemcpyDeviceToHost));
        PCP_CUDA_OK(cudaMemcpy(h_mean, L.mean, cout * 4, cudaMemcpyDeviceToHost));
        PCP_CUDA_OK(cudaMemcpy(h_inv, L.inv, cout * 4, cudaMemcpyDeviceToHost));
        PCP_CUDA_OK(cudaMemcpy(h_dw, dWg, (size_t)128 * cout * 4, cudaMemcpyDeviceToHost));
        PCP_CUDA_OK(cudaMemcpy(h_dgamma, dgam, cout * 4, cudaMemcpyDeviceToHost));
        PCP_CUDA_OK(cudaMemcpy(h_dbeta, dbet, cout * 4, cudaMemcpyDeviceToHost));
        PCP_CUDA_OK(cudaMemcpy(h_da, dAo, rows * 512, cudaMemcpyDeviceToHost));
        return PCP_OK;
    };
    rc = run();
    g_tc = TcCtx();
    g_sync = SyncCtx();
    PCP_RELEASE(cudaFree(arena));
    return rc;
}

extern "C" int pcp_model_set_bn_allreduce(pcp_model *m, pcp_allreduce_fn fn, void *user)
{
    PCP_REQUIRE(m, PCP_ERR_INVALID, "pcp_model_set_bn_allreduce: null model");
    PCP_TRY(ensure_train_buffers(m));
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    ts->sync_fn = fn; ts->sync_user = user;
    return PCP_OK;
}

extern "C" int pcp_model_grad_buffer(pcp_model *m, float **d_grads, size_t *count)
{
    PCP_REQUIRE(m && d_grads && count, PCP_ERR_INVALID, "pcp_model_grad_buffer: null argument");
    PCP_TRY(ensure_train_buffers(m));
    *d_grads = m->d_grads;
    *count = m->n_params;
    return PCP_OK;
}

extern "C" int pcp_model_adam_step(pcp_model *m, const pcp_train_config *cfg, int64_t step, pcp_stream stream)
{
    PCP_REQUIRE(m && cfg && step >= 0, PCP_ERR_INVALID, "pcp_model_adam_step: bad arguments");
    PCP_TRY(ensure_train_buffers(m));
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    // ExponentialDecay(staircase=False): lr = lr0 * rate^(step/decay_steps)  (pointnet_train.py:310-315)
    const double lr = (double)cfg->lr0 * pow((double)cfg->decay_rate, (double)step / (double)cfg->decay_steps);
    const double t = (double)(step + 1);
    const double alpha = lr * sqrt(1.0 - pow((double)cfg->beta2, t)) / (1.0 - pow((double)cfg->beta1, t));
    const size_t n4 = m->n_params / 4;
    adam_kernel<<<sm_count() * 8, 256, 0, as_stream(stream)>>>(reinterpret_cast<float4 *>(m->d_params),
                                                               reinterpret_cast<const float4 *>(m->d_grads),
                                                               reinterpret_cast<float4 *>(m->d_adam_m),
                                                               reinterpret_cast<float4 *>(m->d_adam_v), ts->d_mask, n4,
                                                               (float)alpha, cfg->beta1, cfg->beta2, cfg->eps);
    PCP_CHECK_LAUNCH();
    m->prepared = false;   // inference images are stale until the next prepare
    m->dev_newer = true;
    return PCP_OK;
}

extern "C" int pcp_model_reset_optimizer(pcp_model *m)
{
    PCP_REQUIRE(m, PCP_ERR_INVALID, "pcp_model_reset_optimizer: null model");
    PCP_TRY(ensure_train_buffers(m));
    PCP_CUDA_OK(cudaMemset(m->d_adam_m, 0, m->n_params * sizeof(float)));
    PCP_CUDA_OK(cudaMemset(m->d_adam_v, 0, m->n_params * sizeof(float)));
    return PCP_OK;
}

extern "C" int pcp_model_get_grad(pcp_model *m, const char *name, float *h_data, size_t count)
{
    PCP_REQUIRE(m && name && h_data, PCP_ERR_INVALID, "pcp_model_get_grad: null argument");
    auto it = m->index.find(name);
    PCP_REQUIRE(it != m->index.end(), PCP_ERR_NOTFOUND, "pcp_model_get_grad: unknown tensor '%s'", name);
    const TensorSpec &t = m->tensors[it->second];
    PCP_REQUIRE(t.count == count, PCP_ERR_INVALID, "pcp_model_get_grad: '%s' has %zu elements, got %zu", name, t.count, count);
    PCP_TRY(ensure_train_buffers(m));
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpy(h_data, m->d_grads + t.offset, count * sizeof(float), cudaMemcpyDeviceToHost));
    return PCP_OK;
}

extern "C" int pcp_model_loss_backward(pcp_model *m, const float *d_x, int b, int64_t n, const int32_t *d_y_cls,
                                       const int32_t *d_y_seg, const float *d_y_se3, const pcp_train_config *cfg,
                                       int global_batch, const float *d_mask1, const float *d_mask2, float *h_metrics,
                                       pcp_stream stream)
{
    PCP_REQUIRE(m && d_x && cfg && b >= 1 && n >= 1 && global_batch >= b, PCP_ERR_INVALID, "pcp_model_loss_backward: bad arguments");
    PCP_REQUIRE((int64_t)b * n < ((int64_t)1 << 23), PCP_ERR_UNSUPPORTED, "pcp_model_loss_backward: b*n must stay below 8.3 M points per shard");
    PCP_REQUIRE(!cfg->use_dropout_masks || m->desc.dropout_rate <= 0.f || (d_mask1 && d_mask2), PCP_ERR_INVALID,
                "pcp_model_loss_backward: use_dropout_masks set but masks are null");
    PCP_REQUIRE(m->desc.dropout_rate < 1.f, PCP_ERR_INVALID, "pcp_model_loss_backward: dropout rate must be < 1");
    PCP_TRY(ensure_train_buffers(m));
    // the device copy of the parameters is the master during training
    if (!m->dev_newer && !m->prepared)
        PCP_CUDA_OK(cudaMemcpy(m->d_params, m->h_params.data(), m->n_params * sizeof(float), cudaMemcpyHostToDevice));
    return train_step_impl(m, d_x, b, n, d_y_cls, d_y_seg, d_y_se3, cfg, global_batch, d_mask1, d_mask2, h_metrics, as_stream(stream));
}

// Inference-mode forward + the compiled losses / metrics (keras Model.evaluate / validation pass of fit,
// pointnet_train.py:221-228): no gradients, BN on moving statistics, no dropout.
extern "C" int pcp_model_evaluate(pcp_model *m, const float *d_x, int b, int64_t n, const int32_t *d_y_cls, const int32_t *d_y_seg,
                                  const float *d_y_se3, const pcp_train_config *cfg, float *h_metrics, pcp_stream stream)
{
    PCP_REQUIRE(m && d_x && cfg && h_metrics && b >= 1 && n >= 1, PCP_ERR_INVALID, "pcp_model_evaluate: bad arguments");
    PCP_TRY(ensure_train_buffers(m));
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    cudaStream_t st = as_stream(stream);
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    const int64_t M = (int64_t)b * n;
    const size_t need = (size_t)b * C * 4 + (size_t)M * P * 4 + (size_t)b * 36 + 64 + 1024;
    if (ts->eval_bytes < need) {
        PCP_RELEASE(cudaFree(ts->eval_buf));
        ts->eval_buf = nullptr; ts->eval_bytes = 0;
        PCP_CUDA_OK(cudaMalloc(&ts->eval_buf, need));
        ts->eval_bytes = need;
    }
    double *met = reinterpret_cast<double *>(ts->eval_buf);
    float *cls_logits = reinterpret_cast<float *>(ts->eval_buf + 256);
    float *R = cls_logits + (((size_t)b * C + 63) & ~(size_t)63);
    float *seg_logits = R + (((size_t)b * 9 + 63) & ~(size_t)63);
    const bool want_seg = d_y_seg != nullptr;
    PCP_CUDA_OK(cudaMemsetAsync(met, 0, 64, st));
    // Keras' test_step runs the same forward as training, so the feature T-Net's add_loss is part of val_loss too: the
    // inference forward adds it to met[3] wherever it has R64 in hand
    m->eval_reg_acc = (!m->desc.vanilla && m->desc.regularize_feature_transform) ? met + 3 : nullptr;
    const int frc = pcp_model_forward(m, d_x, b, n, PCP_HEAD_CLS | (want_seg ? PCP_HEAD_SEG : 0u), nullptr, nullptr, R, cls_logits,
                                      want_seg ? seg_logits : nullptr, stream);
    m->eval_reg_acc = nullptr;
    if (frc != PCP_OK) return frc;
    if (d_y_cls) {
        PCP_TRY(softmax_ce(cls_logits, b, C, d_y_cls, 0.f, nullptr, nullptr, met + 0, met + 4, st));
    }
    if (want_seg) {
        PCP_TRY(softmax_ce(seg_logits, M, P, d_y_seg, 0.f, nullptr, nullptr, met + 1, met + 5, st));
    }
    if (d_y_se3) {
        mse_kernel<<<grid_for((int64_t)b * 9, 256, 1), 256, 0, st>>>(R, d_y_se3, (int64_t)b * 9, 0.f, nullptr, met + 2);
        PCP_CHECK_LAUNCH();
    }
    if (!m->desc.vanilla && m->desc.regularize_input_transform) {
        tnet_reg_kernel<<<b, 32, 2 * 9 * sizeof(float), st>>>(R, 3, nullptr, met + 3);
        PCP_CHECK_LAUNCH();
    }
    double hm[8];
    PCP_CUDA_OK(cudaMemcpyAsync(hm, met, 64, cudaMemcpyDeviceToHost, st));
    PCP_CUDA_OK(cudaStreamSynchronize(st));
    const double cls_loss = d_y_cls ? hm[0] / b : 0.0, seg_loss = want_seg ? hm[1] / ((double)b * n) : 0.0;
    const double se3 = d_y_se3 ? hm[2] / ((double)b * 9) : 0.0, reg = hm[3];
    h_metrics[1] = (float)cls_loss; h_metrics[2] = (float)seg_loss; h_metrics[3] = (float)se3;
    h_metrics[4] = d_y_cls ? (float)(hm[4] / b) : 0.f;
    h_metrics[5] = want_seg ? (float)(hm[5] / ((double)b * n)) : 0.f;
    h_metrics[6] = (float)sqrt(se3);
    h_metrics[7] = (float)reg;
    h_metrics[0] = (float)(cfg->w_cls * cls_loss + cfg->w_seg * seg_loss + cfg->w_rot * se3 + reg);
    return PCP_OK;
}

extern "C" int pcp_model_get_train_output(pcp_model *m, int which, float *h_out, size_t count)
{
    PCP_REQUIRE(m && h_out && m->train_ws, PCP_ERR_INVALID, "pcp_model_get_train_output: no training step has run");
    TrainState *ts = reinterpret_cast<TrainState *>(m->train_ws);
    PCP_REQUIRE(ts->arena, PCP_ERR_STATE, "pcp_model_get_train_output: no training step has run");
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    const float *src = nullptr; size_t have = 0;
    switch (which) {
    case 0: src = ts->cls_prob; have = (size_t)ts->b * C; break;
    case 1: src = ts->have_seg ? ts->seg_prob : nullptr; have = (size_t)ts->b * ts->n * P; break;
    case 2: src = ts->R; have = (size_t)ts->b * 9; break;
    case 3: src = m->desc.vanilla ? nullptr : ts->R64; have = (size_t)ts->b * 4096; break;
    case 4: src = ts->g3; have = (size_t)ts->b * 1024; break;
    default: break;
    }
    PCP_REQUIRE(src && have == count, PCP_ERR_INVALID, "pcp_model_get_train_output: output %d unavailable or count mismatch (%zu vs %zu)", which, have, count);
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpy(h_out, src, count * sizeof(float), cudaMemcpyDeviceToHost));
    return PCP_OK;
}
