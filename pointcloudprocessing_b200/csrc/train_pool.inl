// train_pool.inl — the three max-pooled 128 -> 1024 layers of the training step without their (rows x 1024) tensors.
// Included by train.cu (inside namespace pcp, after the BN kernels and the mm_* wrappers).  bf16 precision mode only.
//
// Reference semantics: ConvLayer(1024) -> BN(batch statistics) -> ReLU -> tf.reduce_max over the points
// (pointnet/PointNet.py:425-429, :242-248) and the gradient of that chain.  The explicit path (layer_fwd_train + pool_fwd /
// pool_bwd + bn_bwd + layer_bwd_mm) writes Z = A.W (1 GB at 256 x 1024 points), re-reads it for the statistics, the pool,
// the tie count, the pool gradient, both BN backward passes and the two gradient GEMMs: ~13 GB of HBM traffic per layer.
// None of it is needed:
//   forward   Z is a linear image of the 128-wide input A, so its batch statistics follow from the Gram matrix of A:
//               sum_i z_ij = (1^T A) w_j ,  sum_i z_ij^2 = w_j^T (A^T A) w_j            (A^T A on tcgen05, 128 x 128)
//             and BN+ReLU are monotone per channel, so max_i relu(bn(z_ij)) = relu(bn(max_i z_ij)) for sc_j >= 0 (min for
//             sc_j < 0 — folded away: the GEMM multiplies by W diag(sign gamma), so the pooled row is always the arg max
//             of z' = sign(gamma) z, and z = sign(gamma) z' exactly): one GEMM whose epilogue keeps only the per-cloud
//             column maximum and its row (tc_train.cu tc_pool_extrema).
//   backward  dY is non-zero only at the selected row of each (cloud, channel), and BN backward adds an affine function
//             of Z per channel:  dZ_ij = sc_j dY_ij + c0_j + c1_j z_ij , c1 = -sc inv s2 / M , c0 = -sc s1 / M - c1 mean.
//             With Z = A W:   dW = A^T dZ = (A^T A) W diag(c1) + (A^T 1) c0^T + sum_c S_cj A[arg_cj, :]
//                             dA = dZ W^T = A (W diag(c1) W^T) + 1 (W c0)^T + scatter_c,j S_cj W[:, j] into row arg_cj
//             i.e. one 128 x 128 GEMM over A plus gathers / scatters of b x 1024 rows.
// Ties: tf.reduce_max splits the gradient evenly among tied rows; this path gives it to the first tied row.  For ties
// between duplicated points (clouds padded by repetition, PointCloudSet.py:459-467) every weight gradient is identical,
// because duplicated rows carry identical activations at every layer; for accidental bit-equal maxima of distinct points
// both choices are valid sub-gradients.  Operands are the bf16-rounded A and W the forward GEMM multiplies.

__device__ __forceinline__ float bf16_round(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// asum[k] += sum over rows of bf16(A[r, k])          (cin == 128: 32 float4 columns x 8 row lanes, two rows in flight)
__global__ void __launch_bounds__(256)
pool_colsum_kernel(const float *__restrict__ A, int64_t rows, double *__restrict__ parts)
{
    double *asum = parts + (size_t)blockIdx.x * 128;
    const int c4 = threadIdx.x & 31, lane = threadIdx.x >> 5;
    const float4 *A4 = reinterpret_cast<const float4 *>(A);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    auto add = [&](const float4 v) { acc.x += bf16_round(v.x); acc.y += bf16_round(v.y); acc.z += bf16_round(v.z); acc.w += bf16_round(v.w); };
    const int64_t step = (int64_t)gridDim.x * 8;
    int64_t r = (int64_t)blockIdx.x * 8 + lane;
    for (; r + step < rows; r += 2 * step) {
        const float4 v0 = __ldg(A4 + r * 32 + c4), v1 = __ldg(A4 + (r + step) * 32 + c4);
        add(v0); add(v1);
    }
    for (; r < rows; r += step) add(__ldg(A4 + r * 32 + c4));
    // the eight row lanes of the block first (shared memory), then the block's 128 partial sums (reduce_parts_kernel adds the
    // blocks up: 1024 atomics per block on the same 128 words serialised in the L2 for 0.17 ms per layer, 128 still for 25 us)
    __shared__ float4 part[8][32];
    part[lane][c4] = acc;
    __syncthreads();
    if (lane == 0) {
        double sx = 0, sy = 0, sz = 0, sw = 0;
#pragma unroll
        for (int l = 0; l < 8; ++l) { const float4 v = part[l][c4]; sx += v.x; sy += v.y; sz += v.z; sw += v.w; }
        asum[c4 * 4] = sx; asum[c4 * 4 + 1] = sy; asum[c4 * 4 + 2] = sz; asum[c4 * 4 + 3] = sw;
    }
}

// column sums / sums of squares of Z = A.W from the Gram matrix: one warp per output channel, eight channels per CTA.
// The Gram matrix (64 KB) is staged in shared memory once per CTA: as global loads behind one another the 512 steps of a lane
// were a chain of L2 round trips (49 us per layer; more loads in flight only traded them for register pressure).
constexpr int kStatsWarps = 8;
constexpr size_t kStatsSmem = 128 * 128 * sizeof(float) + kStatsWarps * 128 * sizeof(double);
__global__ void __launch_bounds__(kStatsWarps * 32)
pool_implicit_stats_kernel(const float *__restrict__ gram, const double *__restrict__ asum, const float *__restrict__ W, int cout,
                           double *__restrict__ sum, double *__restrict__ sumsq)
{
    extern __shared__ __align__(16) unsigned char stats_smem[];
    float *gs = reinterpret_cast<float *>(stats_smem);                                    // [128][128]
    double (*ws)[128] = reinterpret_cast<double (*)[128]>(stats_smem + 128 * 128 * sizeof(float));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j = blockIdx.x * kStatsWarps + warp;
    for (int i = threadIdx.x; i < 128 * 32; i += kStatsWarps * 32)
        reinterpret_cast<float4 *>(gs)[i] = __ldg(reinterpret_cast<const float4 *>(gram) + i);
    if (j < cout)
        for (int k = lane; k < 128; k += 32) ws[warp][k] = (double)bf16_round(__ldg(W + (size_t)k * cout + j));
    __syncthreads();
    if (j >= cout) return;
    double s = 0, qd = 0;
    for (int i = 0; i < 4; ++i) {
        const int k = lane + 32 * i;
        double t0 = 0, t1 = 0, t2 = 0, t3 = 0;                 // G symmetric: gs[l][k] is conflict-free over the lanes
#pragma unroll 4
        for (int l = 0; l < 128; l += 4) {
            t0 = fma((double)gs[l * 128 + k], ws[warp][l], t0); t1 = fma((double)gs[(l + 1) * 128 + k], ws[warp][l + 1], t1);
            t2 = fma((double)gs[(l + 2) * 128 + k], ws[warp][l + 2], t2); t3 = fma((double)gs[(l + 3) * 128 + k], ws[warp][l + 3], t3);
        }
        const double t = (t0 + t1) + (t2 + t3);
        qd += ws[warp][k] * t;
        s += ws[warp][k] * asum[k];
    }
    for (int o = 16; o; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); qd += __shfl_xor_sync(0xffffffffu, qd, o); }
    if (lane == 0) { sum[j] = s; sumsq[j] = qd; }
}

// pooled output and the record the backward pass needs
__global__ void __launch_bounds__(256)
pool_select_kernel(const unsigned long long *__restrict__ emax, const float *__restrict__ gamma, BnP p, int64_t total, int cout,
                   float *__restrict__ g, float *__restrict__ zsel, int *__restrict__ arg)
{   // emax: maxima of z' = sign(gamma) z (tc_pool_extrema); the selected extremum of z is sign(gamma) max z'
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int j = (int)(i % cout);
        const float sc = __ldg(p.sc + j);
        const unsigned long long key = emax[i];
        const float zs = __uint_as_float(tc_key_float_bits((uint32_t)(key >> 32)));
        const float z = __ldg(gamma + j) >= 0.f ? zs : -zs;
        zsel[i] = z;
        arg[i] = (int)(0xffffffffu - (uint32_t)key);
        g[i] = fmaxf(bn_pre(z, __ldg(p.mean + j), sc, __ldg(p.beta + j)), 0.f);
    }
}

// BN backward sums over the b selected rows of each channel and S = sc * dY
// (32 channels x 8 cloud lanes per CTA; the clouds are split over gridDim.y CTAs that add into the zeroed sums — as one CTA per 32
//  channels the 32 dependent trips of three loads each took 100 us per layer)
__global__ void __launch_bounds__(256)
pool_bwd_coef_kernel(const float *__restrict__ g, const float *__restrict__ dg, const float *__restrict__ zsel, BnP p, int b, int cout,
                     double *__restrict__ s1o, double *__restrict__ s2o, float *__restrict__ S)
{
    __shared__ double sh1[8][32], sh2[8][32];
    const int jl = threadIdx.x & 31, cl = threadIdx.x >> 5;
    const int j = blockIdx.x * 32 + jl;
    const bool ok = j < cout;
    const float mean = ok ? p.mean[j] : 0.f, inv = ok ? p.inv[j] : 0.f, sc = ok ? p.sc[j] : 0.f;
    double s1 = 0, s2 = 0;
    if (ok)
        for (int c = blockIdx.y * 8 + cl; c < b; c += 8 * gridDim.y) {
            const size_t i = (size_t)c * cout + j;
            const float dy = g[i] > 0.f ? dg[i] : 0.f;
            s1 += dy;
            s2 += (double)dy * (double)((zsel[i] - mean) * inv);
            S[i] = sc * dy;
        }
    sh1[cl][jl] = s1; sh2[cl][jl] = s2;
    __syncthreads();
    if (cl == 0 && ok) {
        for (int l = 1; l < 8; ++l) { s1 += sh1[l][jl]; s2 += sh2[l][jl]; }
        atomicAdd(s1o + j, s1); atomicAdd(s2o + j, s2);
    }
}

// c1 = -sc inv s2 / M , c0 = -sc s1 / M - c1 mean from the (all-reduced) sums; zero for a frozen BatchNorm
__global__ void __launch_bounds__(128)
pool_coef_kernel(const double *__restrict__ s1, const double *__restrict__ s2, BnP p, int cout, double rows, int batch, float *__restrict__ c0,
                 float *__restrict__ c1)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= cout) return;
    if (batch) {
        const double sc = p.sc[j], k1 = -sc * (double)p.inv[j] * s2[j] / rows;
        c1[j] = (float)k1;
        c0[j] = (float)(-sc * s1[j] / rows - k1 * (double)p.mean[j]);
    } else {
        c1[j] = 0.f; c0[j] = 0.f;
    }
}

// Wt (cout x cin) = bf16(W (cin x cout))^T
__global__ void __launch_bounds__(256)
pool_round_transpose_kernel(const float *__restrict__ W, int cin, int cout, float *__restrict__ Wt)
{
    __shared__ float tile[32][33];
    const int j0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int r = ty; r < 32; r += 8) tile[r][tx] = (k0 + r < cin && j0 + tx < cout) ? bf16_round(__ldg(W + (size_t)(k0 + r) * cout + j0 + tx)) : 0.f;
    __syncthreads();
    for (int r = ty; r < 32; r += 8)
        if (j0 + r < cout && k0 + tx < cin) Wt[(size_t)(j0 + r) * cin + k0 + tx] = tile[tx][r];
}

// dW[k, j] += c1_j (G Wr)[k, j] + asum[k] c0_j + sum_c S[c, j] bf16(A[c n + arg[c, j], k])       (cin == 128; 8 channels per CTA)
__global__ void __launch_bounds__(128)
pool_dw_kernel(const float *__restrict__ gram, const double *__restrict__ asum, const float *__restrict__ Wt, const float *__restrict__ c0,
               const float *__restrict__ c1, const float *__restrict__ S, const int *__restrict__ arg, const float *__restrict__ A, int b, int64_t n,
               int cout, int batch, float *__restrict__ dW)
{
    const int k = threadIdx.x, j0 = blockIdx.x * 8;
    float acc[8];
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) acc[jj] = 0.f;
    // the last row of CTAs (blockIdx.y == gridDim.y - 1) takes the Gram term, the others the clouds: behind one another in the same
    // CTA the two dependent-load chains added up to 67 us per layer
    const int ycl = (int)gridDim.y - 1;
    if ((int)blockIdx.y == ycl) {
        if (batch) {
#pragma unroll 4
            for (int l = 0; l < 128; ++l) {
                const float gv = __ldg(gram + l * 128 + k);
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) acc[jj] = fmaf(gv, __ldg(Wt + (size_t)(j0 + jj) * 128 + l), acc[jj]);
            }
            const float as = (float)asum[k];
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) acc[jj] = acc[jj] * __ldg(c1 + j0 + jj) + as * __ldg(c0 + j0 + jj);
        }
    } else {
        // four clouds per trip: 32 (S, arg) pairs, then 32 row gathers in flight — the kernel is a chain of dependent L2 / HBM
        // round trips (ncu: 17.7 of 22 stalled warps on the long scoreboard, 25 MB of DRAM traffic in 66 us)
        for (int c = blockIdx.y; c < b; c += 4 * ycl) {
            float s[32], a[32];
            int r[32];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int cu = c + u * ycl;
                const bool in = cu < b;
                const size_t i = (size_t)(in ? cu : c) * cout + j0;
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) { s[u * 8 + jj] = in ? __ldg(S + i + jj) : 0.f; r[u * 8 + jj] = __ldg(arg + i + jj); }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int cu = c + u * ycl < b ? c + u * ycl : c;
#pragma unroll
                for (int jj = 0; jj < 8; ++jj) a[u * 8 + jj] = __ldg(A + ((int64_t)cu * n + r[u * 8 + jj]) * 128 + k);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int jj = 0; jj < 8; ++jj)
                    if (s[u * 8 + jj] != 0.f) acc[jj] = fmaf(s[u * 8 + jj], bf16_round(a[u * 8 + jj]), acc[jj]);
        }
    }
    // two 16-byte reductions per thread (cout and j0 are multiples of 8: aligned)
    float4 *dst = reinterpret_cast<float4 *>(dW + (size_t)k * cout + j0);
    atomicAdd(dst, make_float4(acc[0], acc[1], acc[2], acc[3]));
    atomicAdd(dst + 1, make_float4(acc[4], acc[5], acc[6], acc[7]));
}

// Q[k, l] = sum_j Wr[k, j] c1_j Wr[l, j] ; q[k] = sum_j Wr[k, j] c0_j             (cin == 128; one CTA of 8 x 128 threads per k)
// Thread (grp, l) sums the grp-th eighth of the channels for column l (128 threads walking all 1024 channels, a handful of loads
// in flight each, took 47 us per layer); the eight partial sums are added in group order.
__global__ void __launch_bounds__(1024)
pool_q_kernel(const float *__restrict__ Wt, const float *__restrict__ c0, const float *__restrict__ c1, int cout, float *__restrict__ Q,
              float *__restrict__ q)
{
    extern __shared__ float sk[];                 // cout floats: Wr[k, j] c1_j
    __shared__ float part[32];
    __shared__ float red[8][128];
    const int k = blockIdx.x, t = threadIdx.x, l = t & 127, grp = t >> 7;
    float qk = 0.f;
    for (int j = t; j < cout; j += 1024) {
        const float w = __ldg(Wt + (size_t)j * 128 + k);
        sk[j] = w * __ldg(c1 + j);
        qk = fmaf(w, __ldg(c0 + j), qk);
    }
    for (int o = 16; o; o >>= 1) qk += __shfl_xor_sync(0xffffffffu, qk, o);
    if ((t & 31) == 0) part[t >> 5] = qk;
    __syncthreads();
    const int per = (cout + 7) / 8, j0 = grp * per, j1 = min(cout, j0 + per);
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;          // independent chains: the loop is load-latency bound
    int j = j0;
#pragma unroll 2
    for (; j + 3 < j1; j += 4) {
        a0 = fmaf(sk[j], __ldg(Wt + (size_t)j * 128 + l), a0); a1 = fmaf(sk[j + 1], __ldg(Wt + (size_t)(j + 1) * 128 + l), a1);
        a2 = fmaf(sk[j + 2], __ldg(Wt + (size_t)(j + 2) * 128 + l), a2); a3 = fmaf(sk[j + 3], __ldg(Wt + (size_t)(j + 3) * 128 + l), a3);
    }
    for (; j < j1; ++j) a0 = fmaf(sk[j], __ldg(Wt + (size_t)j * 128 + l), a0);
    red[grp][l] = (a0 + a1) + (a2 + a3);
    __syncthreads();
    if (t < 128) {
        float v = 0.f;
#pragma unroll
        for (int gg = 0; gg < 8; ++gg) v += red[gg][t];
        Q[k * 128 + t] = v;
    }
    if (t == 0) {
        float v = 0.f;
        for (int w = 0; w < 32; ++w) v += part[w];
        q[k] = v;
    }
}

// dA[c n + arg[c, j], k] += S[c, j] Wr[k, j]                                                       (cin == 128)
__global__ void __launch_bounds__(128)
pool_scatter_kernel(const float *__restrict__ S, const int *__restrict__ arg, const float *__restrict__ Wt, int64_t n, int cout, int jchunk,
                    float *__restrict__ dA)
{
    const int k = threadIdx.x, c = blockIdx.y, j0 = blockIdx.x * jchunk, j1 = min(cout, j0 + jchunk);
    for (int j = j0; j < j1; j += 8) {               // eight channels' loads in flight: one dependent S -> arg -> Wt chain per channel was 46 us per layer
        float s[8], w[8];
        int r[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const bool in = j + e < j1;
            const size_t i = (size_t)c * cout + j + e;
            s[e] = in ? __ldg(S + i) : 0.f;
            r[e] = in ? __ldg(arg + i) : 0;
            w[e] = in ? __ldg(Wt + (size_t)(j + e) * 128 + k) : 0.f;
        }
#pragma unroll
        for (int e = 0; e < 8; ++e)
            if (s[e] != 0.f) atomicAdd(dA + ((int64_t)c * n + r[e]) * 128 + k, s[e] * w[e]);
    }
}

struct PoolLayer {                      // everything one pooled layer's implicit forward / backward touches
    const float *W = nullptr;           // (128 x cout) kernel
    const float *gamma = nullptr, *beta_in = nullptr;
    float *mov_mean = nullptr, *mov_var = nullptr;
    float *mean = nullptr, *inv = nullptr, *sc = nullptr, *beta = nullptr;     // per-step BN record (cout each)
    double *red = nullptr;              // 2 x cout
    int b = 0; int64_t n = 0; int cout = 0;
    bool batch = false, trainable = false;
    PoolBufs pb;
    const __nv_bfloat16 *A16 = nullptr; // optional: the layer's input already rounded to bf16 (same values the GEMMs round to)
    BnP bnp() const { return BnP{mean, sc, beta, inv}; }
};

static bool pool_implicit_ok(int cin, int cout) { return g_tc.on && cin == 128 && cout % 128 == 0 && gemm_b_image_bytes(cin, cout) <= g_tc.img_bytes; }

// g (b x cout) = max over the cloud's points of relu(bn(A . W)); BN record + moving statistics updated like bn_prepare
static int pooled_forward(const PoolLayer &L, const float *A, float *g, cudaStream_t st)
{
    const int64_t rows = (int64_t)L.b * L.n, total = (int64_t)L.b * L.cout;
    PCP_CUDA_OK(cudaMemsetAsync(L.pb.ext, 0, (size_t)total * 8, st));
    const uint8_t *wimg = nullptr;
    PCP_TRY(tc_image(L.W, 128, L.cout, L.cout, false, &wimg, st, L.gamma));          // W . diag(sign(gamma))
    PCP_TRY(tc_pool_extrema(A, L.b, L.n, wimg, L.cout, L.pb.ext, st));
    if (L.batch) {
        // the Gram matrix feeds the layer's batch statistics, i.e. the FORWARD: it has to be the same bits on every run
        PCP_TRY(tc_gram_det(A, 128, rows, 128, L.pb.gram, L.pb.gram_parts, tc_gram_scratch_floats(128), st));
        const int np = (int)std::min<int64_t>((rows + 7) / 8, (int64_t)sm_count() * 4);
        pool_colsum_kernel<<<np, 256, 0, st>>>(A, rows, L.pb.colparts);
        PCP_CHECK_LAUNCH();
        reduce_parts_kernel<<<128, 256, 0, st>>>(L.pb.colparts, np, 128, L.pb.asum, nullptr, nullptr, 0, nullptr);
        PCP_CHECK_LAUNCH();
        // statistics use the Gram matrix / column sums of the CONCATENATED batch (sync-BN); dW below keeps this rank's
        const float *gram_s = L.pb.gram;
        const double *asum_s = L.pb.asum;
        if (g_sync.fn) {
            PCP_CUDA_OK(cudaMemcpyAsync(L.pb.gram_g, L.pb.gram, 128 * 128 * 4, cudaMemcpyDeviceToDevice, st));
            PCP_CUDA_OK(cudaMemcpyAsync(L.pb.asum_g, L.pb.asum, 128 * 8, cudaMemcpyDeviceToDevice, st));
            PCP_TRY(sync_sum(L.pb.gram_g, 128 * 128, 0, st));
            PCP_TRY(sync_sum(L.pb.asum_g, 128, 1, st));
            gram_s = L.pb.gram_g; asum_s = L.pb.asum_g;
        }
        static bool stats_attr[64] = {};
        int dev = 0;
        PCP_CUDA_OK(cudaGetDevice(&dev));
        if (dev >= 0 && dev < 64 && !stats_attr[dev]) {
            PCP_CUDA_OK(cudaFuncSetAttribute(pool_implicit_stats_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kStatsSmem));
            stats_attr[dev] = true;
        }
        pool_implicit_stats_kernel<<<(L.cout + kStatsWarps - 1) / kStatsWarps, kStatsWarps * 32, kStatsSmem, st>>>(gram_s, asum_s, L.W, L.cout, L.red,
                                                                                                              L.red + L.cout);
        PCP_CHECK_LAUNCH();
    }
    bn_finalize_kernel<<<L.cout, 256, 0, st>>>(L.red, 1, (double)rows * g_sync.scale, L.cout, L.batch ? 1 : 0, L.gamma, L.beta_in, L.mov_mean,
                                                        L.mov_var, L.mean, L.inv, L.sc, L.beta, kBnEpsT, kBnMomentum);
    PCP_CHECK_LAUNCH();
    pool_select_kernel<<<grid_for(total), 256, 0, st>>>(L.pb.ext, L.gamma, L.bnp(), total, L.cout, g, L.pb.zsel, L.pb.arg);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// dg (b x cout) -> dW (+=), dgamma / dbeta (=), dA (rows x 128, =); A must be the forward input again
static int pooled_backward(const PoolLayer &L, const float *A, const float *g, const float *dg, float *dW, float *dgamma, float *dbeta, float *dA,
                           cudaStream_t st)
{
    const int64_t rows = (int64_t)L.b * L.n;
    PCP_CUDA_OK(cudaMemsetAsync(L.red, 0, (size_t)L.cout * 16, st));
    pool_bwd_coef_kernel<<<dim3((L.cout + 31) / 32, (unsigned)std::max(1, std::min(32, (L.b + 7) / 8))), 256, 0, st>>>(g, dg, L.pb.zsel, L.bnp(), L.b, L.cout,
                                                                                                                  L.red, L.red + L.cout, L.pb.S);
    PCP_CHECK_LAUNCH();
    if (L.batch && L.trainable && dgamma && dbeta) {          // this rank's sums: the gradient bucket is summed over the ranks later
        store_bn_grads_kernel<<<(L.cout + 127) / 128, 128, 0, st>>>(L.red, L.red + L.cout, L.cout, dgamma, dbeta);
        PCP_CHECK_LAUNCH();
    }
    if (L.batch) PCP_TRY(sync_sum(L.red, (size_t)L.cout * 2, 1, st));
    pool_coef_kernel<<<(L.cout + 127) / 128, 128, 0, st>>>(L.red, L.red + L.cout, L.bnp(), L.cout, (double)rows * g_sync.scale, L.batch ? 1 : 0, L.pb.c0,
                                                         L.pb.c1);
    PCP_CHECK_LAUNCH();
    pool_round_transpose_kernel<<<dim3((L.cout + 31) / 32, 4), 256, 0, st>>>(L.W, 128, L.cout, L.pb.Wt);
    PCP_CHECK_LAUNCH();
    if (L.trainable && dW) {
        pool_dw_kernel<<<dim3(L.cout / 8, std::min(L.b, 16) + 1), 128, 0, st>>>(L.pb.gram, L.pb.asum, L.pb.Wt, L.pb.c0, L.pb.c1, L.pb.S, L.pb.arg, A, L.b, L.n, L.cout,
                                                  L.batch ? 1 : 0, dW);
        PCP_CHECK_LAUNCH();
    }
    if (dA) {
        if (L.batch) {
            pool_q_kernel<<<128, 1024, (size_t)L.cout * 4, st>>>(L.pb.Wt, L.pb.c0, L.pb.c1, L.cout, L.pb.Q, L.pb.q);
            PCP_CHECK_LAUNCH();
            PCP_TRY(mm_nn(A, rows, 128, L.pb.Q, 128, 128, dA, L.pb.q, st, 1, 0, L.A16));
        } else {
            PCP_CUDA_OK(cudaMemsetAsync(dA, 0, (size_t)rows * 512, st));
        }
        pool_scatter_kernel<<<dim3((L.cout + 63) / 64, L.b), 128, 0, st>>>(L.pb.S, L.pb.arg, L.pb.Wt, L.n, L.cout, 64, dA);
        PCP_CHECK_LAUNCH();
    }
    return PCP_OK;
}
