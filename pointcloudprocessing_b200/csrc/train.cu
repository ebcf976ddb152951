// train.cu — training step of pointnet_train.py:310-351: losses, backward, fused Adam.
#include <cmath>
#include <cstring>

#include "gemm_simt.cuh"
#include "model.cuh"

namespace pcp {

// Fused keras Adam over the flat parameter buffer (pointnet_train.py:317-319):
//   m = b1*m + (1-b1)*g ; v = b2*v + (1-b2)*g^2 ; w -= alpha * m / (sqrt(v) + eps),
//   alpha = lr * sqrt(1-b2^t) / (1-b1^t).  28 B/parameter of HBM traffic (read w,g,m,v; write w,m,v).
// `mask` holds one byte per 4-float group: 0 => frozen / BN statistic, skipped.
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

struct TrainState {
    unsigned char *d_mask = nullptr;
    bool mask_valid = false;
    bool trainable_seen[4] = {false, false, false, false};
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

}  // namespace pcp

using namespace pcp;

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
    (void)m; (void)d_x; (void)b; (void)n; (void)d_y_cls; (void)d_y_seg; (void)d_y_se3; (void)cfg; (void)global_batch;
    (void)d_mask1; (void)d_mask2; (void)h_metrics; (void)stream;
    set_error("pcp_model_loss_backward: training backward is not built in this revision");
    return PCP_ERR_UNSUPPORTED;
}
