// model.cu — pcp_model lifecycle, tensor table, inference BN fold, and the fp32 forward pass
// (PointNet.call, pointnet/PointNet.py:197-292) expressed as a sequence of fused GEMM launches.
#include <cmath>
#include <cstring>

#include "gemm_simt.cuh"
#include "model.cuh"
#include "tc_gemm.cuh"
#include "tc_train.cuh"

namespace pcp {

static const float kBnEps = 1e-3f;   // keras BatchNormalization default epsilon

static void add_tensor(pcp_model *m, const std::string &name, size_t count, int group, bool is_stat, int *slot)
{
    TensorSpec t;
    t.name = name; t.offset = m->n_params; t.count = count; t.group = group; t.is_stat = is_stat;
    m->n_params += (count + 3) & ~(size_t)3;     // keep every tensor 16-byte aligned in the flat buffer
    m->index[name] = (int)m->tensors.size();
    if (slot) *slot = (int)m->tensors.size();
    m->tensors.push_back(t);
}

static void add_layer(pcp_model *m, int id, const std::string &name, int cin, int cout, bool has_bn, int group)
{
    LayerSpec &L = m->layers[id];
    L.name = name; L.cin = cin; L.cout = cout; L.has_bn = has_bn; L.present = true; L.group = group;
    add_tensor(m, name + "/kernel", (size_t)cin * cout, group, false, &L.t_kernel);
    if (has_bn) {
        add_tensor(m, name + "/bn/gamma", cout, group, false, &L.t_gamma);
        add_tensor(m, name + "/bn/beta", cout, group, false, &L.t_beta);
        add_tensor(m, name + "/bn/moving_mean", cout, group, true, &L.t_mean);
        add_tensor(m, name + "/bn/moving_variance", cout, group, true, &L.t_var);
    } else {
        add_tensor(m, name + "/bias", cout, group, false, &L.t_bias);
    }
}

static void add_tnet(pcp_model *m, int first, const std::string &tn, int k, int group, int *t_w, int *t_b)
{
    const int w[5] = {64, 128, 1024, 512, 256};          // layer_widths, PointNet.py:384
    add_layer(m, first + 0, tn + "_convolution_layer_1_convolution_layer", k, w[0], true, group);
    add_layer(m, first + 1, tn + "_convolution_layer_2_convolution_layer", w[0], w[1], true, group);
    add_layer(m, first + 2, tn + "_convolution_layer_3_convolution_layer", w[1], w[2], true, group);
    add_layer(m, first + 3, tn + "_dense_layer_1_dense_layer", w[2], w[3], true, group);
    add_layer(m, first + 4, tn + "_dense_layer_2_dense_layer", w[3], w[4], true, group);
    add_tensor(m, tn + "/w", (size_t)256 * k * k, group, false, t_w);
    add_tensor(m, tn + "/b", (size_t)k * k, group, false, t_b);
}

static void build_spec(pcp_model *m)
{
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    const bool full = !m->desc.vanilla;
    if (full) add_tnet(m, L_IT_C1, "input_transform", 3, PCP_GROUP_INPUT_TRANSFORM, &m->t_it_w, &m->t_it_b);
    add_layer(m, L_S1_1, "s1_l1_64_convolution_layer", 3, 64, true, PCP_GROUP_SHARED);
    add_layer(m, L_S1_2, "s1_l2_64_convolution_layer", 64, 64, true, PCP_GROUP_SHARED);
    if (full) add_tnet(m, L_FT_C1, "feature_transform", 64, PCP_GROUP_SHARED, &m->t_ft_w, &m->t_ft_b);
    add_layer(m, L_S2_1, "s2_l1_64_convolution_layer", 64, 64, true, PCP_GROUP_SHARED);
    add_layer(m, L_S2_2, "s2_l2_128_convolution_layer", 64, 128, true, PCP_GROUP_SHARED);
    add_layer(m, L_S2_3, "s2_l3_1024_convolution_layer", 128, 1024, true, PCP_GROUP_SHARED);
    add_layer(m, L_CLS1, "s3_l1_512_dense_layer", 1024, 512, true, PCP_GROUP_CLS_HEAD);
    add_layer(m, L_CLS2, "s3_l2_256_dense_layer", 512, 256, true, PCP_GROUP_CLS_HEAD);
    add_layer(m, L_CLS3, "output_dense_layer", 256, C, false, PCP_GROUP_CLS_HEAD);
    add_layer(m, L_SEG1, "seg_l1_512_convolution_layer", 1088, 512, true, PCP_GROUP_SEG_HEAD);
    add_layer(m, L_SEG2, "seg_l2_256_convolution_layer", 512, 256, true, PCP_GROUP_SEG_HEAD);
    add_layer(m, L_SEG3, "seg_l3_128_convolution_layer", 256, 128, true, PCP_GROUP_SEG_HEAD);
    add_layer(m, L_SEG4, "seg_l4_128_convolution_layer", 128, 128, true, PCP_GROUP_SEG_HEAD);
    add_layer(m, L_SEG5, "seg_l5_output_convolution_layer", 128, P, false, PCP_GROUP_SEG_HEAD);
    // folded image layout
    size_t off = 0;
    auto take = [&](size_t cnt) { size_t o = off; off += (cnt + 3) & ~(size_t)3; return o; };
    for (int i = 0; i < L_COUNT; ++i) {
        LayerSpec &L = m->layers[i];
        if (!L.present) continue;
        L.fold_w = take((size_t)L.cin * L.cout);
        L.fold_t = take(L.cout);
    }
    if (full) {
        m->fold_it_w = take(256 * 9); m->fold_it_t = take(9);
        m->fold_ft_w = take(256 * 4096); m->fold_ft_t = take(4096);
    }
    m->n_fold = off;
}

static void init_defaults(pcp_model *m)
{
    // Keras build-time state: BN gamma=1, beta=0, mean=0, var=1 (PointNet.py:528); TNet b = identity (:416).
    // Kernels are left at zero: the host side (pointnet/PointNet.py mirror) draws Glorot-uniform values.
    for (auto &t : m->tensors) {
        const std::string &nm = t.name;
        float fill = 0.f;
        if (nm.size() > 9 && nm.compare(nm.size() - 9, 9, "/bn/gamma") == 0) fill = 1.f;
        if (nm.size() > 19 && nm.compare(nm.size() - 19, 19, "/bn/moving_variance") == 0) fill = 1.f;
        if (fill != 0.f) for (size_t i = 0; i < t.count; ++i) m->h_params[t.offset + i] = fill;
    }
    auto eye = [&](int ti, int k) {
        if (ti < 0) return;
        for (int i = 0; i < k; ++i) m->h_params[m->tensors[ti].offset + (size_t)i * k + i] = 1.f;
    };
    eye(m->t_it_b, 3);
    eye(m->t_ft_b, 64);
}

static int ws_alloc(Workspace &ws, float **p, size_t count)
{
    void *q = nullptr;
    PCP_CUDA_OK(cudaMalloc(&q, (count ? count : 1) * sizeof(float)));
    ws.allocs.push_back(q);
    *p = reinterpret_cast<float *>(q);
    return PCP_OK;
}

static void ws_free(Workspace &ws)
{
    for (void *p : ws.allocs) PCP_RELEASE(cudaFree(p));
    if (ws.pin_x) PCP_RELEASE(cudaFreeHost(ws.pin_x));
    if (ws.pin_out) PCP_RELEASE(cudaFreeHost(ws.pin_out));
    if (ws.copy_stream) PCP_RELEASE(cudaStreamDestroy(ws.copy_stream));
    if (ws.ev_start) PCP_RELEASE(cudaEventDestroy(ws.ev_start));
    for (cudaEvent_t e : ws.ev_copied) PCP_RELEASE(cudaEventDestroy(e));
    ws = Workspace();
}

__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float *__restrict__ z, float *__restrict__ p, int64_t rows, int cols)
{
    for (int64_t r = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
        const float *zr = z + r * cols;
        float mx = zr[0];
        for (int c = 1; c < cols; ++c) mx = fmaxf(mx, zr[c]);
        float s = 0.f;
        for (int c = 0; c < cols; ++c) s += expf(zr[c] - mx);
        const float inv = 1.0f / s;
        for (int c = 0; c < cols; ++c) p[r * cols + c] = expf(zr[c] - mx) * inv;
    }
}

__global__ void eye3_kernel(float *R, int b)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < b) {
#pragma unroll
        for (int j = 0; j < 9; ++j) R[i * 9 + j] = (j % 4 == 0) ? 1.f : 0.f;
    }
}

int model_sync_host(pcp_model *m)
{
    if (m->dev_newer) {
        PCP_CUDA_OK(cudaDeviceSynchronize());
        PCP_CUDA_OK(cudaMemcpy(m->h_params.data(), m->d_params, m->n_params * sizeof(float), cudaMemcpyDeviceToHost));
        m->dev_newer = false;
    }
    return PCP_OK;
}

int softmax_rows(const float *d_logits, float *d_prob, int64_t rows, int cols, cudaStream_t st)
{
    if (rows <= 0) return PCP_OK;
    int64_t blocks = (rows + 255) / 256;
    if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
    softmax_rows_kernel<<<(unsigned)blocks, 256, 0, st>>>(d_logits, d_prob, rows, cols);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// ---- PCP_PRECISION_FP32_TC: the fp32 forward with its layer products on the bf16 tensor cores (hi / lo split) ----
struct Tc3Images {
    uint8_t *split[L_COUNT] = {};        // [hi(W'); lo(W'); hi(W')] image of every layer with cin % 64 == 0 (tc_gemm a_split3)
    uint8_t *pool_hi[L_COUNT] = {}, *pool_lo[L_COUNT] = {};   // the max-pooled 128 -> 1024 layers: hi / lo images for tc_pool_max3
    uint8_t *seg1_local = nullptr;       // seg_l1 rows [0, 64): the X_64 part of the hoisted 1088 -> 512 layer
    std::vector<void *> owned;
};

static void tc3_destroy(pcp_model *m)
{
    if (!m->tc3) return;
    Tc3Images *t = reinterpret_cast<Tc3Images *>(m->tc3);
    for (void *p : t->owned) PCP_RELEASE(cudaFree(p));
    delete t;
    m->tc3 = nullptr;
}

static int tc3_upload(Tc3Images *t, const std::vector<uint16_t> &img, uint8_t **out)
{
    void *d = nullptr;
    PCP_CUDA_OK(cudaMalloc(&d, img.size() * 2));
    PCP_CUDA_OK(cudaMemcpy(d, img.data(), img.size() * 2, cudaMemcpyHostToDevice));
    t->owned.push_back(d);
    *out = reinterpret_cast<uint8_t *>(d);
    return PCP_OK;
}

static int tc3_prepare(pcp_model *m, const std::vector<float> &fold)
{
    tc3_destroy(m);
    Tc3Images *t = new Tc3Images();
    m->tc3 = t;
    std::vector<uint16_t> img;
    for (int i = 0; i < L_COUNT; ++i) {
        const LayerSpec &L = m->layers[i];
        if (!L.present || L.cin % 64 != 0 || !(L.cout <= 128 || L.cout % 128 == 0)) continue;
        const float *W = fold.data() + L.fold_w;
        if (i == L_IT_C3 || i == L_FT_C3 || i == L_S2_3) {
            std::vector<float> hi((size_t)L.cin * L.cout), lo(hi.size());
            for (size_t k = 0; k < hi.size(); ++k) {
                uint32_t u; memcpy(&u, &W[k], 4);
                uint32_t r = u + 0x7FFFu + ((u >> 16) & 1u);            // bf16 round-to-nearest-even (finite weights)
                r &= 0xFFFF0000u;
                memcpy(&hi[k], &r, 4);
                lo[k] = W[k] - hi[k];
            }
            pack_gemm_b_image(hi.data(), L.cin, L.cout, img);
            PCP_TRY(tc3_upload(t, img, &t->pool_hi[i]));
            pack_gemm_b_image(lo.data(), L.cin, L.cout, img);
            PCP_TRY(tc3_upload(t, img, &t->pool_lo[i]));
        } else if (i == L_SEG1) {
            pack_gemm_b_image_split3(W, 64, L.cout, img);                // rows [0, 64) of the (1088 x 512) kernel
            PCP_TRY(tc3_upload(t, img, &t->seg1_local));
        } else {
            pack_gemm_b_image_split3(W, L.cin, L.cout, img);
            PCP_TRY(tc3_upload(t, img, &t->split[i]));
        }
    }
    return PCP_OK;
}

// C[g] (M x N) = act(A[g] . W' + bias[g]) through tc_gemm with the hi / lo split
static int tc3_gemm(const float *A, int64_t rows, int cin, const uint8_t *img, int cout, float *Cc, const float *bias, bool relu, cudaStream_t st,
                    int groups = 1, int64_t bias_stride = 0)
{
    TcGemmParams p;
    const int64_t rpg = rows / groups;
    p.A = A; p.a_f32 = 1; p.a_split3 = 1; p.lda = cin; p.a_group_stride = rpg * cin; p.Bimg = img; p.bias = bias; p.bias_group_stride = bias_stride;
    p.C32 = Cc; p.ldc = cout; p.c_group_stride = rpg * cout; p.M = (int)rpg; p.N = cout; p.K = 3 * cin; p.groups = groups; p.relu = relu ? 1 : 0;
    return tc_gemm(p, st);
}

// C = act(A . Wfold + t): one ConvLayer / DenseLayer in inference form (PointNet.py:554-566, :642-654)
static int layer_fwd(pcp_model *m, int id, const float *A, int64_t M, float *C, bool relu, cudaStream_t st,
                     float *colmax = nullptr, int groups = 1, int64_t rows_per_group = 0)
{
    const LayerSpec &L = m->layers[id];
    if (m->desc.precision == PCP_PRECISION_FP32_TC && m->tc3) {
        const Tc3Images *t3 = reinterpret_cast<const Tc3Images *>(m->tc3);
        if (colmax && t3->pool_hi[id] && L.cin == 128)
            return tc_pool_max3(A, groups, rows_per_group, t3->pool_hi[id], t3->pool_lo[id], m->d_fold + L.fold_t, L.cout, colmax, st);
        if (!colmax && t3->split[id] && (((uintptr_t)A) & 15u) == 0 && M < 0x7fffffff && (M + 127) / 128 <= 65535)
            return tc3_gemm(A, M, L.cin, t3->split[id], L.cout, C, m->d_fold + L.fold_t, relu, st);
    }
    GemmArgs g;
    g.A = A; g.lda = L.cin; g.B = m->d_fold + L.fold_w; g.ldb = L.cout; g.C = C; g.ldc = L.cout;
    g.bias = m->d_fold + L.fold_t; g.N = L.cout; g.K = L.cin; g.relu = relu ? 1 : 0;
    if (colmax) {
        g.groups = groups; g.M = (int)rows_per_group; g.strideA = rows_per_group * L.cin;
        g.strideC = rows_per_group * L.cout; g.colmax = colmax; g.strideColmax = L.cout;
    } else {
        g.M = (int)M;
    }
    return gemm_f32(g, st);
}

// TNet tail: pooled (bc,1024) -> dense 512 -> 256 -> (256 x K^2) + b  (PointNet.py:431-442)
static int tnet_tail(pcp_model *m, int first, size_t fold_w, size_t fold_t, int k, const float *pooled, int bc, float *out,
                     cudaStream_t st)
{
    Workspace &ws = m->ws;
    PCP_TRY(layer_fwd(m, first + 3, pooled, bc, ws.d512, true, st));
    PCP_TRY(layer_fwd(m, first + 4, ws.d512, bc, ws.d256, true, st));
    GemmArgs g;
    g.A = ws.d256; g.lda = 256; g.B = m->d_fold + fold_w; g.ldb = k * k; g.C = out; g.ldc = k * k;
    g.bias = m->d_fold + fold_t; g.M = bc; g.N = k * k; g.K = 256;
    return gemm_f32(g, st);
}


// Segmentation head (PointNet.py:268-290) from X_64 (bc*n, 64) and the global features (bc, 1024).
// concat([X_64, tile(global)]) . W1  ==  X_64 . W1[0:64] + (global . W1[64:1088]) — the second term is
// constant per cloud and becomes a per-cloud bias vector (the tile/concat tensors never exist).
static int seg_from_x64(pcp_model *m, const float *X64, const float *g3, int bc, int64_t n, float *d_seg, float *d_seg_logits,
                        cudaStream_t st)
{
    Workspace &ws = m->ws;
    const int P = m->desc.num_parts;
    const int64_t M = (int64_t)bc * n;
    const LayerSpec &S1 = m->layers[L_SEG1];
    GemmArgs gv;
    gv.A = g3; gv.lda = 1024; gv.B = m->d_fold + S1.fold_w + (size_t)64 * 512; gv.ldb = 512;
    gv.C = ws.gvec; gv.ldc = 512; gv.bias = m->d_fold + S1.fold_t; gv.M = bc; gv.N = 512; gv.K = 1024;
    PCP_TRY(gemm_f32(gv, st));
    const Tc3Images *t3 = (m->desc.precision == PCP_PRECISION_FP32_TC) ? reinterpret_cast<const Tc3Images *>(m->tc3) : nullptr;
    if (t3 && t3->seg1_local && (int64_t)bc * ((n + 127) / 128) <= 65535) {
        PCP_TRY(tc3_gemm(X64, M, 64, t3->seg1_local, 512, ws.s512, ws.gvec, true, st, bc, 512));
    } else {
        GemmArgs g;
        g.A = X64; g.lda = 64; g.strideA = n * 64; g.B = m->d_fold + S1.fold_w; g.ldb = 512; g.strideB = 0;
        g.C = ws.s512; g.ldc = 512; g.strideC = n * 512; g.bias = ws.gvec; g.strideBias = 512;
        g.M = (int)n; g.N = 512; g.K = 64; g.groups = bc; g.relu = 1;
        PCP_TRY(gemm_f32(g, st));
    }
    PCP_TRY(layer_fwd(m, L_SEG2, ws.s512, M, ws.s256, true, st));
    PCP_TRY(layer_fwd(m, L_SEG3, ws.s256, M, ws.s128a, true, st));
    PCP_TRY(layer_fwd(m, L_SEG4, ws.s128a, M, ws.s128b, true, st));
    float *logits = d_seg_logits ? d_seg_logits : ws.seglog;
    PCP_TRY(layer_fwd(m, L_SEG5, ws.s128b, M, logits, false, st));
    if (d_seg) PCP_TRY(softmax_rows(logits, d_seg, M, P, st));
    return PCP_OK;
}

static int ensure_point_ws(pcp_model *m);
static int model_cls_head_chunk_impl(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st);
static int model_cls_head_chunk(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st)
{
    return model_cls_head_chunk_impl(m, g3, b, d_cls, d_cls_logits, st);
}

int model_tnet_tail(pcp_model *m, int which, const float *pooled, int bc, float *out, cudaStream_t st)
{
    return which == 0 ? tnet_tail(m, L_IT_C1, m->fold_it_w, m->fold_it_t, 3, pooled, bc, out, st)
                      : tnet_tail(m, L_FT_C1, m->fold_ft_w, m->fold_ft_t, 64, pooled, bc, out, st);
}

int model_eye3(float *d_R, int b, cudaStream_t st)
{
    eye3_kernel<<<(b + 255) / 256, 256, 0, st>>>(d_R, b);
    PCP_CHECK_LAUNCH();
    return PCP_OK;
}

// classification head (PointNet.py:252-263): dense 512 -> 256 -> C (+bias) -> softmax
int model_cls_head(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st)
{
    return model_cls_head_chunk(m, g3, b, d_cls, d_cls_logits, st);
}

static int model_cls_head_chunk_impl(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st)
{
    Workspace &ws = m->ws;
    const int C = m->desc.num_classes;
    PCP_TRY(layer_fwd(m, L_CLS1, g3, b, ws.d512, true, st));
    PCP_TRY(layer_fwd(m, L_CLS2, ws.d512, b, ws.d256, true, st));
    float *logits = d_cls_logits ? d_cls_logits : ws.clslog;
    PCP_TRY(layer_fwd(m, L_CLS3, ws.d256, b, logits, false, st));
    if (d_cls) PCP_TRY(softmax_rows(logits, d_cls, b, C, st));
    return PCP_OK;
}

// Segmentation head for the bf16 trunk: X_64 is rebuilt in fp32 from the raw points (4 288 MAC/point)
// and the head runs on the fp32 GEMM path.  R / R64 null => vanilla.
int model_seg_head_fp32(pcp_model *m, const float *d_x, int b, int64_t n, const float *d_R, const float *d_R64, const float *g3,
                        float *d_seg, float *d_seg_logits, cudaStream_t st)
{
    PCP_TRY(ensure_point_ws(m));
    Workspace &ws = m->ws;
    const int P = m->desc.num_parts;
    for (int c0 = 0; c0 < b; c0 += ws.cap_b) {
        const int bc = std::min(ws.cap_b, b - c0);
        const int64_t M = (int64_t)bc * n;
        PCP_TRY(pcp_normalize_f32(d_x + (int64_t)c0 * n * 3, bc, n, ws.pcn, nullptr, nullptr, st));
        const float *X3 = ws.pcn;
        if (d_R) {
            GemmArgs g;
            g.A = ws.pcn; g.lda = 3; g.strideA = n * 3; g.B = d_R + (int64_t)c0 * 9; g.ldb = 3; g.strideB = 9;
            g.C = ws.xr; g.ldc = 3; g.strideC = n * 3; g.M = (int)n; g.N = 3; g.K = 3; g.groups = bc;
            PCP_TRY(gemm_f32(g, st));
            X3 = ws.xr;
        }
        PCP_TRY(layer_fwd(m, L_S1_1, X3, M, ws.h64a, true, st));
        PCP_TRY(layer_fwd(m, L_S1_2, ws.h64a, M, ws.h64b, true, st));
        const float *X64 = ws.h64b;
        if (d_R64) {
            GemmArgs g;
            g.A = ws.h64b; g.lda = 64; g.strideA = n * 64; g.B = d_R64 + (int64_t)c0 * 4096; g.ldb = 64; g.strideB = 4096;
            g.C = ws.x64; g.ldc = 64; g.strideC = n * 64; g.M = (int)n; g.N = 64; g.K = 64; g.groups = bc;
            PCP_TRY(gemm_f32(g, st));
            X64 = ws.x64;
        }
        PCP_TRY(seg_from_x64(m, X64, g3 + (int64_t)c0 * 1024, bc, n, d_seg ? d_seg + (int64_t)c0 * n * P : nullptr,
                             d_seg_logits ? d_seg_logits + (int64_t)c0 * n * P : nullptr, st));
    }
    return PCP_OK;
}

int model_forward_fp32(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg,
                       float *d_R, float *d_cls_logits, float *d_seg_logits, cudaStream_t st)
{
    Workspace &ws = m->ws;
    const bool full = !m->desc.vanilla;
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    for (int c0 = 0; c0 < b; c0 += ws.cap_b) {
        const int bc = std::min(ws.cap_b, b - c0);
        const int64_t M = (int64_t)bc * n;
        const float *x = d_x + (int64_t)c0 * n * 3;
        float *g1 = ws.g1 + (int64_t)c0 * 1024, *g2 = ws.g2 + (int64_t)c0 * 1024, *g3 = ws.g3 + (int64_t)c0 * 1024;
        PCP_TRY(pcp_normalize_f32(x, bc, n, ws.pcn, nullptr, nullptr, st));                       // :202
        const float *X3 = ws.pcn;
        float *Rout = d_R ? d_R + (int64_t)c0 * 9 : ws.R;
        if (full) {
            PCP_CUDA_OK(cudaMemsetAsync(g1, 0, (size_t)bc * 1024 * sizeof(float), st));
            PCP_TRY(layer_fwd(m, L_IT_C1, ws.pcn, M, ws.h64a, true, st));                          // TNet.call :423
            PCP_TRY(layer_fwd(m, L_IT_C2, ws.h64a, M, ws.h128, true, st));                         // :424
            PCP_TRY(layer_fwd(m, L_IT_C3, ws.h128, M, nullptr, true, st, g1, bc, n));              // :425 + reduce_max :429
            PCP_TRY(tnet_tail(m, L_IT_C1, m->fold_it_w, m->fold_it_t, 3, g1, bc, Rout, st));       // :431-442
            GemmArgs g;                                                                            // X = pc @ R  :207
            g.A = ws.pcn; g.lda = 3; g.strideA = n * 3; g.B = Rout; g.ldb = 3; g.strideB = 9;
            g.C = ws.xr; g.ldc = 3; g.strideC = n * 3; g.M = (int)n; g.N = 3; g.K = 3; g.groups = bc;
            PCP_TRY(gemm_f32(g, st));
            X3 = ws.xr;
        } else if (d_R) {
            eye3_kernel<<<(bc + 255) / 256, 256, 0, st>>>(Rout, bc);                                // :211
            PCP_CHECK_LAUNCH();
        }
        PCP_TRY(layer_fwd(m, L_S1_1, X3, M, ws.h64a, true, st));                                   // :217
        PCP_TRY(layer_fwd(m, L_S1_2, ws.h64a, M, ws.h64b, true, st));                              // :220
        const float *X64 = ws.h64b;
        if (full) {
            PCP_CUDA_OK(cudaMemsetAsync(g2, 0, (size_t)bc * 1024 * sizeof(float), st));
            PCP_TRY(layer_fwd(m, L_FT_C1, ws.h64b, M, ws.h64a, true, st));
            PCP_TRY(layer_fwd(m, L_FT_C2, ws.h64a, M, ws.h128, true, st));
            PCP_TRY(layer_fwd(m, L_FT_C3, ws.h128, M, nullptr, true, st, g2, bc, n));
            PCP_TRY(tnet_tail(m, L_FT_C1, m->fold_ft_w, m->fold_ft_t, 64, g2, bc, ws.R64, st));     // :227
            if (m->eval_reg_acc) PCP_TRY(tnet_reg_accumulate(ws.R64, bc, 64, m->eval_reg_acc, st));
            GemmArgs g;                                                                            // X_64 = X @ R_64 :228
            g.A = ws.h64b; g.lda = 64; g.strideA = n * 64; g.B = ws.R64; g.ldb = 64; g.strideB = 4096;
            g.C = ws.x64; g.ldc = 64; g.strideC = n * 64; g.M = (int)n; g.N = 64; g.K = 64; g.groups = bc;
            PCP_TRY(gemm_f32(g, st));
            X64 = ws.x64;
        }
        PCP_CUDA_OK(cudaMemsetAsync(g3, 0, (size_t)bc * 1024 * sizeof(float), st));
        PCP_TRY(layer_fwd(m, L_S2_1, X64, M, ws.h64a, true, st));                                  // :236
        PCP_TRY(layer_fwd(m, L_S2_2, ws.h64a, M, ws.h128, true, st));                              // :239
        PCP_TRY(layer_fwd(m, L_S2_3, ws.h128, M, nullptr, true, st, g3, bc, n));                   // :242 + :248

        if (heads & PCP_HEAD_CLS)                                                                  // :252-263
            PCP_TRY(model_cls_head_chunk(m, g3, bc, d_cls ? d_cls + (int64_t)c0 * C : nullptr,
                                         d_cls_logits ? d_cls_logits + (int64_t)c0 * C : nullptr, st));
        if (heads & PCP_HEAD_SEG)
            PCP_TRY(seg_from_x64(m, X64, g3, bc, n, d_seg ? d_seg + (int64_t)c0 * n * P : nullptr,
                                 d_seg_logits ? d_seg_logits + (int64_t)c0 * n * P : nullptr, st));
    }
    return PCP_OK;
}

int model_forward_bf16(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg,
                       float *d_R, float *d_cls_logits, float *d_seg_logits, cudaStream_t st);
int model_forward_bf16_begin(pcp_model *m, int b, cudaStream_t st);
int model_forward_bf16_stage1(pcp_model *m, const float *d_x_chunk, int c0, int bc, int64_t n, cudaStream_t st);
int model_forward_bf16_finish(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg, float *d_R,
                              float *d_cls_logits, float *d_seg_logits, cudaStream_t st);
int model_prepare_bf16(pcp_model *m);
void model_destroy_bf16(pcp_model *m);

}  // namespace pcp

using namespace pcp;

namespace pcp {
static int ensure_point_ws(pcp_model *m)
{
    Workspace &ws = m->ws;
    if (ws.pcn) return PCP_OK;
    const int P = m->desc.num_parts;
    const size_t Mc = (size_t)ws.cap_b * (size_t)ws.cap_n;
    PCP_TRY(ws_alloc(ws, &ws.pcn, Mc * 3)); PCP_TRY(ws_alloc(ws, &ws.xr, Mc * 3));
    PCP_TRY(ws_alloc(ws, &ws.h64a, Mc * 64)); PCP_TRY(ws_alloc(ws, &ws.h64b, Mc * 64));
    PCP_TRY(ws_alloc(ws, &ws.x64, Mc * 64)); PCP_TRY(ws_alloc(ws, &ws.h128, Mc * 128));
    PCP_TRY(ws_alloc(ws, &ws.s512, Mc * 512)); PCP_TRY(ws_alloc(ws, &ws.s256, Mc * 256));
    PCP_TRY(ws_alloc(ws, &ws.s128a, Mc * 128)); PCP_TRY(ws_alloc(ws, &ws.s128b, Mc * 128));
    PCP_TRY(ws_alloc(ws, &ws.seglog, Mc * P));
    return PCP_OK;
}
}  // namespace pcp

extern "C" int pcp_model_create(const pcp_model_desc *desc, pcp_model **out)
{
    PCP_REQUIRE(desc && out, PCP_ERR_INVALID, "pcp_model_create: null argument");
    PCP_REQUIRE(desc->num_classes >= 1 && desc->num_parts >= 1, PCP_ERR_INVALID, "pcp_model_create: widths must be >= 1");
    PCP_REQUIRE(desc->precision == PCP_PRECISION_FP32 || desc->precision == PCP_PRECISION_BF16 || desc->precision == PCP_PRECISION_FP32_TC, PCP_ERR_INVALID,
                "pcp_model_create: unknown precision %d", desc->precision);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error("pcp_model_create: no CUDA device (%s) — this library has no CPU fallback", cudaGetErrorString(e));
        return PCP_ERR_CUDA;
    }
    pcp_model *m = new pcp_model();
    m->desc = *desc;
    build_spec(m);
    m->h_params.assign(m->n_params, 0.f);
    init_defaults(m);
    if (cudaMalloc(&m->d_params, m->n_params * sizeof(float)) != cudaSuccess ||
        cudaMalloc(&m->d_fold, m->n_fold * sizeof(float)) != cudaSuccess) {
        set_error("pcp_model_create: cudaMalloc failed");
        delete m;
        return PCP_ERR_NOMEM;
    }
    *out = m;
    return PCP_OK;
}

extern "C" int pcp_model_destroy(pcp_model *m)
{
    if (!m) return PCP_OK;
    ws_free(m->ws);
    model_destroy_bf16(m);
    tc3_destroy(m);
    train_destroy(m);
    PCP_RELEASE(cudaFree(m->d_params)); PCP_RELEASE(cudaFree(m->d_fold));
    PCP_RELEASE(cudaFree(m->d_grads)); PCP_RELEASE(cudaFree(m->d_adam_m)); PCP_RELEASE(cudaFree(m->d_adam_v));
    delete m;
    return PCP_OK;
}

extern "C" int pcp_model_num_tensors(const pcp_model *m) { return m ? (int)m->tensors.size() : 0; }

extern "C" int pcp_model_tensor_info(const pcp_model *m, int index, char *name_buf, size_t name_buf_len, size_t *count,
                                     int *trainable_group)
{
    PCP_REQUIRE(m && index >= 0 && index < (int)m->tensors.size(), PCP_ERR_INVALID, "pcp_model_tensor_info: bad index");
    const TensorSpec &t = m->tensors[index];
    if (name_buf && name_buf_len) { strncpy(name_buf, t.name.c_str(), name_buf_len - 1); name_buf[name_buf_len - 1] = 0; }
    if (count) *count = t.count;
    if (trainable_group) *trainable_group = t.is_stat ? -1 : t.group;
    return PCP_OK;
}

extern "C" int pcp_model_set_tensor(pcp_model *m, const char *name, const float *h_data, size_t count)
{
    PCP_REQUIRE(m && name && h_data, PCP_ERR_INVALID, "pcp_model_set_tensor: null argument");
    auto it = m->index.find(name);
    PCP_REQUIRE(it != m->index.end(), PCP_ERR_NOTFOUND, "pcp_model_set_tensor: unknown tensor '%s'", name);
    const TensorSpec &t = m->tensors[it->second];
    PCP_REQUIRE(t.count == count, PCP_ERR_INVALID, "pcp_model_set_tensor: '%s' has %zu elements, got %zu", name, t.count, count);
    PCP_TRY(model_sync_host(m));
    memcpy(m->h_params.data() + t.offset, h_data, count * sizeof(float));
    m->prepared = false;
    return PCP_OK;
}

extern "C" int pcp_model_get_tensor(pcp_model *m, const char *name, float *h_data, size_t count)
{
    PCP_REQUIRE(m && name && h_data, PCP_ERR_INVALID, "pcp_model_get_tensor: null argument");
    auto it = m->index.find(name);
    PCP_REQUIRE(it != m->index.end(), PCP_ERR_NOTFOUND, "pcp_model_get_tensor: unknown tensor '%s'", name);
    const TensorSpec &t = m->tensors[it->second];
    PCP_REQUIRE(t.count == count, PCP_ERR_INVALID, "pcp_model_get_tensor: '%s' has %zu elements, got %zu", name, t.count, count);
    PCP_TRY(model_sync_host(m));
    memcpy(h_data, m->h_params.data() + t.offset, count * sizeof(float));
    return PCP_OK;
}

extern "C" int pcp_model_prepare(pcp_model *m)
{
    PCP_REQUIRE(m, PCP_ERR_INVALID, "pcp_model_prepare: null model");
    PCP_TRY(model_sync_host(m));
    std::vector<float> fold(m->n_fold, 0.f);
    const float *P = m->h_params.data();
    for (int i = 0; i < L_COUNT; ++i) {
        const LayerSpec &L = m->layers[i];
        if (!L.present) continue;
        const float *W = P + m->tensors[L.t_kernel].offset;
        float *Wf = fold.data() + L.fold_w, *tf = fold.data() + L.fold_t;
        if (L.has_bn) {
            const float *g = P + m->tensors[L.t_gamma].offset, *be = P + m->tensors[L.t_beta].offset;
            const float *mu = P + m->tensors[L.t_mean].offset, *var = P + m->tensors[L.t_var].offset;
            for (int c = 0; c < L.cout; ++c) {
                const float s = g[c] / sqrtf(var[c] + kBnEps);       // inference BN: y = x*s + t
                tf[c] = be[c] - mu[c] * s;
                for (int k = 0; k < L.cin; ++k) Wf[(size_t)k * L.cout + c] = W[(size_t)k * L.cout + c] * s;
            }
        } else {
            const float *bias = P + m->tensors[L.t_bias].offset;
            memcpy(Wf, W, (size_t)L.cin * L.cout * sizeof(float));
            memcpy(tf, bias, L.cout * sizeof(float));
        }
    }
    if (!m->desc.vanilla) {
        memcpy(fold.data() + m->fold_it_w, P + m->tensors[m->t_it_w].offset, 256 * 9 * sizeof(float));
        memcpy(fold.data() + m->fold_it_t, P + m->tensors[m->t_it_b].offset, 9 * sizeof(float));
        memcpy(fold.data() + m->fold_ft_w, P + m->tensors[m->t_ft_w].offset, (size_t)256 * 4096 * sizeof(float));
        memcpy(fold.data() + m->fold_ft_t, P + m->tensors[m->t_ft_b].offset, 4096 * sizeof(float));
    }
    PCP_CUDA_OK(cudaMemcpy(m->d_fold, fold.data(), m->n_fold * sizeof(float), cudaMemcpyHostToDevice));
    PCP_CUDA_OK(cudaMemcpy(m->d_params, m->h_params.data(), m->n_params * sizeof(float), cudaMemcpyHostToDevice));
    if (m->desc.precision == PCP_PRECISION_BF16) PCP_TRY(model_prepare_bf16(m));
    if (m->desc.precision == PCP_PRECISION_FP32_TC) PCP_TRY(tc3_prepare(m, fold));
    m->prepared = true;
    return PCP_OK;
}

extern "C" int pcp_model_reserve(pcp_model *m, int b, int64_t n)
{
    PCP_REQUIRE(m && b >= 1 && n >= 1, PCP_ERR_INVALID, "pcp_model_reserve: bad arguments");
    Workspace &ws = m->ws;
    if (ws.total_b >= b && ws.cap_n == n) return PCP_OK;
    ws_free(ws);
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    const bool fp32 = m->desc.precision != PCP_PRECISION_BF16;
    // fp32 path materialises per-point activations: bound a chunk to ~192k points
    const int64_t max_pts = 196608;
    int cap_b = (int)std::max<int64_t>(1, std::min<int64_t>(b, max_pts / n));
    ws.cap_b = cap_b; ws.cap_n = n; ws.total_b = b;
    if (fp32) PCP_TRY(ensure_point_ws(m));
    PCP_TRY(ws_alloc(ws, &ws.g1, (size_t)b * 1024)); PCP_TRY(ws_alloc(ws, &ws.g2, (size_t)b * 1024));
    PCP_TRY(ws_alloc(ws, &ws.g3, (size_t)b * 1024));
    PCP_TRY(ws_alloc(ws, &ws.d512, (size_t)b * 512)); PCP_TRY(ws_alloc(ws, &ws.d256, (size_t)b * 256));
    PCP_TRY(ws_alloc(ws, &ws.R, (size_t)b * 9)); PCP_TRY(ws_alloc(ws, &ws.R64, (size_t)b * 4096));
    PCP_TRY(ws_alloc(ws, &ws.gvec, (size_t)b * 512)); PCP_TRY(ws_alloc(ws, &ws.clslog, (size_t)b * C));
    return PCP_OK;
}

extern "C" int pcp_model_forward(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg,
                                 float *d_R, float *d_cls_logits, float *d_seg_logits, pcp_stream stream)
{
    PCP_REQUIRE(m && d_x && b >= 1 && n >= 1, PCP_ERR_INVALID, "pcp_model_forward: bad arguments");
    if (!m->prepared) PCP_TRY(pcp_model_prepare(m));      // parameters changed (set_tensor / training step): refold lazily
    PCP_REQUIRE(!(d_cls || d_cls_logits) || (heads & PCP_HEAD_CLS), PCP_ERR_INVALID, "pcp_model_forward: cls output without PCP_HEAD_CLS");
    PCP_REQUIRE(!(d_seg || d_seg_logits) || (heads & PCP_HEAD_SEG), PCP_ERR_INVALID, "pcp_model_forward: seg output without PCP_HEAD_SEG");
    if (m->ws.total_b < b || m->ws.cap_n != n) PCP_TRY(pcp_model_reserve(m, b, n));
    if (m->desc.precision == PCP_PRECISION_BF16)
        return model_forward_bf16(m, d_x, b, n, heads, d_cls, d_seg, d_R, d_cls_logits, d_seg_logits, as_stream(stream));
    return model_forward_fp32(m, d_x, b, n, heads, d_cls, d_seg, d_R, d_cls_logits, d_seg_logits, as_stream(stream));
}

extern "C" int pcp_model_predict_host(pcp_model *m, const float *h_x, int b, int64_t n, uint32_t heads, float *h_cls,
                                      float *h_seg, float *h_R, pcp_stream stream)
{
    PCP_REQUIRE(m && h_x && b >= 1 && n >= 1, PCP_ERR_INVALID, "pcp_model_predict_host: bad arguments");
    if (m->ws.total_b < b || m->ws.cap_n != n) PCP_TRY(pcp_model_reserve(m, b, n));
    Workspace &ws = m->ws;
    const int C = m->desc.num_classes, P = m->desc.num_parts;
    const size_t xb = (size_t)b * n * 3 * sizeof(float);
    const size_t cls_cnt = (h_cls ? (size_t)b * C : 0), seg_cnt = (h_seg ? (size_t)b * n * P : 0), r_cnt = (h_R ? (size_t)b * 9 : 0);
    cudaStream_t st = as_stream(stream);
    if (!ws.stage_x) {
        PCP_TRY(ws_alloc(ws, &ws.stage_x, (size_t)ws.total_b * ws.cap_n * 3));
        PCP_TRY(ws_alloc(ws, &ws.stage_cls, (size_t)ws.total_b * C));
        PCP_TRY(ws_alloc(ws, &ws.stage_seg, (size_t)ws.total_b * ws.cap_n * P));
        PCP_TRY(ws_alloc(ws, &ws.stage_R, (size_t)ws.total_b * 9));
    }
    // The caller's buffers are used directly: pinned memory (cudaMallocHost / torch pin_memory) is DMA'd at full PCIe
    // rate, pageable memory is staged by the driver.  The batch is cut into chunks: a copy stream moves chunk i+1 to the
    // device while the compute stream runs the forward of chunk i, and each chunk's results go back as soon as they exist.
    if (!ws.copy_stream) {
        PCP_CUDA_OK(cudaStreamCreateWithFlags(&ws.copy_stream, cudaStreamNonBlocking));
        PCP_CUDA_OK(cudaEventCreateWithFlags(&ws.ev_start, cudaEventDisableTiming));
    }
    // bf16 model, no per-point output: only the part of the forward that needs nothing but the points (statistics + first chain)
    // runs per chunk, behind the chunk's copy; the T-Net tails, the other two chains and the heads run ONCE on the whole batch
    // (a full forward per chunk pays the dependent-launch latency of the tails and short chain launches three times over).
    // 1/8 of the batch per chunk (measured at 4096 x 1024: 4 / 6 / 8 / 12 / 16 / 24 chunks -> 3.40 / 3.35 / 3.35 / 3.39 / 3.42 / 3.64 ms):
    // the first chain consumes points about as fast as PCIe delivers them, so what stays exposed is the last chunk's chain, while
    // short chain launches lose efficiency.
    static const bool no_staged = getenv("PCP_PREDICT_NO_STAGED") != nullptr;
    if (!no_staged && m->desc.precision == PCP_PRECISION_BF16 && !h_seg) {
        if (!m->prepared) PCP_TRY(pcp_model_prepare(m));
        static const int parts = getenv("PCP_PREDICT_CHUNKS") ? std::max(1, atoi(getenv("PCP_PREDICT_CHUNKS"))) : 8;      // tuning knob
        const int cb = std::max(1, std::max((int)(((int64_t)1 << 17) / n), (b + parts - 1) / parts));
        const int nchunks = (b + cb - 1) / cb;
        while ((int)ws.ev_copied.size() < nchunks) {
            cudaEvent_t e;
            PCP_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ws.ev_copied.push_back(e);
        }
        PCP_CUDA_OK(cudaEventRecord(ws.ev_start, st));
        PCP_CUDA_OK(cudaStreamWaitEvent(ws.copy_stream, ws.ev_start, 0));
        for (int i = 0; i < nchunks; ++i) {
            const int c0 = i * cb, bc = std::min(cb, b - c0);
            PCP_CUDA_OK(cudaMemcpyAsync(ws.stage_x + (size_t)c0 * n * 3, h_x + (size_t)c0 * n * 3, (size_t)bc * n * 3 * sizeof(float),
                                        cudaMemcpyHostToDevice, ws.copy_stream));
            PCP_CUDA_OK(cudaEventRecord(ws.ev_copied[i], ws.copy_stream));
        }
        PCP_TRY(model_forward_bf16_begin(m, b, st));
        for (int i = 0; i < nchunks; ++i) {
            const int c0 = i * cb, bc = std::min(cb, b - c0);
            PCP_CUDA_OK(cudaStreamWaitEvent(st, ws.ev_copied[i], 0));
            PCP_TRY(model_forward_bf16_stage1(m, ws.stage_x + (size_t)c0 * n * 3, c0, bc, n, st));
        }
        PCP_TRY(model_forward_bf16_finish(m, ws.stage_x, b, n, heads, h_cls ? ws.stage_cls : nullptr, nullptr, h_R ? ws.stage_R : nullptr, nullptr,
                                          nullptr, st));
        if (h_cls) PCP_CUDA_OK(cudaMemcpyAsync(h_cls, ws.stage_cls, cls_cnt * sizeof(float), cudaMemcpyDeviceToHost, st));
        if (h_R) PCP_CUDA_OK(cudaMemcpyAsync(h_R, ws.stage_R, r_cnt * sizeof(float), cudaMemcpyDeviceToHost, st));
        PCP_CUDA_OK(cudaStreamSynchronize(st));
        return PCP_OK;
    }
    // Otherwise the batch is cut into chunks that each run a whole forward.
    // Chunk sizes grow geometrically (256 K points, then x3): the first copy is short, so compute starts early, and every
    // later copy hides behind the forward of the chunk before it while the per-forward fixed cost is paid only ~4 times.
    std::vector<int> bounds(1, 0);
    {
        int64_t pts = (int64_t)1 << 18;
        while (bounds.back() < b) {
            int cb = (int)std::max<int64_t>(1, pts / n);
            if (b - bounds.back() - cb < cb) cb = b - bounds.back();      // fold a short tail into this chunk
            bounds.push_back(std::min(b, bounds.back() + cb));
            pts *= 3;
        }
    }
    const int nchunks = (int)bounds.size() - 1;
    while ((int)ws.ev_copied.size() < nchunks) {
        cudaEvent_t e;
        PCP_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        ws.ev_copied.push_back(e);
    }
    // the copy stream may not overwrite the staging buffer before earlier work on the caller's stream has finished
    PCP_CUDA_OK(cudaEventRecord(ws.ev_start, st));
    PCP_CUDA_OK(cudaStreamWaitEvent(ws.copy_stream, ws.ev_start, 0));
    for (int i = 0; i < nchunks; ++i) {
        const int c0 = bounds[i], bc = bounds[i + 1] - c0;
        PCP_CUDA_OK(cudaMemcpyAsync(ws.stage_x + (size_t)c0 * n * 3, h_x + (size_t)c0 * n * 3, (size_t)bc * n * 3 * sizeof(float),
                                    cudaMemcpyHostToDevice, ws.copy_stream));
        PCP_CUDA_OK(cudaEventRecord(ws.ev_copied[i], ws.copy_stream));
    }
    for (int i = 0; i < nchunks; ++i) {
        const int c0 = bounds[i], bc = bounds[i + 1] - c0;
        PCP_CUDA_OK(cudaStreamWaitEvent(st, ws.ev_copied[i], 0));
        PCP_TRY(pcp_model_forward(m, ws.stage_x + (size_t)c0 * n * 3, bc, n, heads, h_cls ? ws.stage_cls + (size_t)c0 * C : nullptr,
                                  h_seg ? ws.stage_seg + (size_t)c0 * n * P : nullptr, h_R ? ws.stage_R + (size_t)c0 * 9 : nullptr,
                                  nullptr, nullptr, stream));
        if (h_seg)       // the only large result: stream it back per chunk
            PCP_CUDA_OK(cudaMemcpyAsync(h_seg + (size_t)c0 * n * P, ws.stage_seg + (size_t)c0 * n * P, (size_t)bc * n * P * sizeof(float),
                                        cudaMemcpyDeviceToHost, st));
    }
    (void)xb; (void)seg_cnt;
    if (h_cls) PCP_CUDA_OK(cudaMemcpyAsync(h_cls, ws.stage_cls, cls_cnt * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (h_R) PCP_CUDA_OK(cudaMemcpyAsync(h_R, ws.stage_R, r_cnt * sizeof(float), cudaMemcpyDeviceToHost, st));
    PCP_CUDA_OK(cudaStreamSynchronize(st));
    return PCP_OK;
}

extern "C" int pcp_model_get_pooled(pcp_model *m, int which, int b, float *h_out)
{
    PCP_REQUIRE(m && h_out && which >= 0 && which <= 2 && b >= 1 && b <= m->ws.total_b, PCP_ERR_INVALID, "pcp_model_get_pooled: bad arguments");
    const float *src = which == 0 ? m->ws.g1 : (which == 1 ? m->ws.g2 : m->ws.g3);
    PCP_CUDA_OK(cudaDeviceSynchronize());
    PCP_CUDA_OK(cudaMemcpy(h_out, src, (size_t)b * 1024 * sizeof(float), cudaMemcpyDeviceToHost));
    return PCP_OK;
}

extern "C" int pcp_model_set_trainable(pcp_model *m, int group, int flag)
{
    PCP_REQUIRE(m && group >= 0 && group < 4, PCP_ERR_INVALID, "pcp_model_set_trainable: bad group %d", group);
    m->trainable[group] = flag != 0;
    return PCP_OK;
}
