// model.cuh — the pcp_model object: flat fp32 parameter buffer keyed by the reference's layer names
// (pointnet/PointNet.py:116-141, :406-416), folded inference images, activation workspace.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "common.cuh"

namespace pcp {

enum LayerId {
    L_IT_C1 = 0, L_IT_C2, L_IT_C3, L_IT_D1, L_IT_D2,       // input_transform TNet (PointNet.py:406-410)
    L_S1_1, L_S1_2,                                         // s1_l1_64, s1_l2_64 (:120-121)
    L_FT_C1, L_FT_C2, L_FT_C3, L_FT_D1, L_FT_D2,            // feature_transform TNet
    L_S2_1, L_S2_2, L_S2_3,                                 // s2_l1_64, s2_l2_128, s2_l3_1024 (:125-127)
    L_CLS1, L_CLS2, L_CLS3,                                 // s3_l1_512, s3_l2_256, output (:130-134)
    L_SEG1, L_SEG2, L_SEG3, L_SEG4, L_SEG5,                 // seg_l1_512 .. seg_l5_output (:137-141)
    L_COUNT
};

struct LayerSpec {
    std::string name;     // e.g. "s1_l1_64_convolution_layer"
    int cin = 0, cout = 0;
    bool has_bn = false;  // else has bias (use_bias = not apply_bn, PointNet.py:540,630)
    bool present = false; // T-Net layers are absent in vanilla mode
    int group = 0;        // PCP_GROUP_*
    // indices into pcp_model::tensors (-1 if absent)
    int t_kernel = -1, t_bias = -1, t_gamma = -1, t_beta = -1, t_mean = -1, t_var = -1;
    // folded inference image (offsets in floats into d_fold): W' = W*s (cin x cout), t (cout)
    size_t fold_w = 0, fold_t = 0;
};

struct TensorSpec {
    std::string name;
    size_t offset = 0;   // floats into the flat parameter buffer
    size_t count = 0;
    int group = 0;
    bool is_stat = false;  // BN moving statistics: not trained by Adam
};

struct Workspace {
    int cap_b = 0;           // clouds per chunk
    int64_t cap_n = 0;
    int total_b = 0;
    float *pcn = nullptr, *xr = nullptr, *h64a = nullptr, *h64b = nullptr, *x64 = nullptr, *h128 = nullptr;
    float *s512 = nullptr, *s256 = nullptr, *s128a = nullptr, *s128b = nullptr, *seglog = nullptr;
    float *g1 = nullptr, *g2 = nullptr, *g3 = nullptr;                 // pooled features (total_b x 1024)
    float *d512 = nullptr, *d256 = nullptr, *R = nullptr, *R64 = nullptr, *gvec = nullptr, *clslog = nullptr;
    float *stage_x = nullptr;   // device staging for predict_host
    float *stage_cls = nullptr, *stage_seg = nullptr, *stage_R = nullptr;
    float *pin_x = nullptr, *pin_out = nullptr;   // pinned host staging
    size_t pin_x_bytes = 0, pin_out_bytes = 0;
    cudaStream_t copy_stream = nullptr;            // predict_host: H2D of chunk i+1 overlaps the forward of chunk i
    cudaEvent_t ev_start = nullptr;
    std::vector<cudaEvent_t> ev_copied;
    std::vector<void *> allocs;
};

}  // namespace pcp

struct pcp_model {
    pcp_model_desc desc;
    std::vector<pcp::TensorSpec> tensors;
    std::map<std::string, int> index;
    pcp::LayerSpec layers[pcp::L_COUNT];
    int t_it_w = -1, t_it_b = -1, t_ft_w = -1, t_ft_b = -1;   // TNet w (256,K^2), b (K,K) (PointNet.py:415-416)
    size_t fold_it_w = 0, fold_it_t = 0, fold_ft_w = 0, fold_ft_t = 0;
    size_t n_params = 0;
    std::vector<float> h_params;
    float *d_params = nullptr;
    float *d_fold = nullptr;
    size_t n_fold = 0;
    bool prepared = false;
    double *eval_reg_acc = nullptr;      // pcp_model_evaluate: device accumulator the inference forward adds the feature T-Net's
                                         // orthogonality regulariser to (per chunk, since R64 only exists chunk by chunk); null otherwise
    bool dev_newer = false;    // d_params was updated on the device (Adam) after the last host sync
    bool trainable[4] = {true, true, true, true};
    pcp::Workspace ws;
    // training state (train.cu)
    float *d_grads = nullptr, *d_adam_m = nullptr, *d_adam_v = nullptr;
    void *train_ws = nullptr;
    // bf16 tensor-core images (chain_tc.cu)
    void *tc = nullptr;
    void *tc_heads = nullptr;   // heads_tc.cu
    void *tc3 = nullptr;        // PCP_PRECISION_FP32_TC: split weight images (model.cu)
};

namespace pcp {
int model_sync_host(pcp_model *m);
int model_forward_fp32(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg,
                       float *d_R, float *d_cls_logits, float *d_seg_logits, cudaStream_t st);
int softmax_rows(const float *d_logits, float *d_prob, int64_t rows, int cols, cudaStream_t st);
int model_tnet_tail(pcp_model *m, int which, const float *pooled, int bc, float *out, cudaStream_t st);
int model_eye3(float *d_R, int b, cudaStream_t st);
int model_cls_head(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st);
int heads_tc_prepare(pcp_model *m, const std::vector<float> &fold);
void heads_tc_destroy(pcp_model *m);
// 1e-3 * sum_b 1/2 ||I - M_b M_b^T||_F^2 of b (k x k) matrices added to *acc (train.cu; PointNet.py:447-451)
int tnet_reg_accumulate(const float *M, int b, int k, double *acc, cudaStream_t st);
int heads_tc_tnet_tail(pcp_model *m, int which, const float *pooled, int b, float *out, cudaStream_t st);
int heads_tc_tnet_tail_fold(pcp_model *m, const float *pooled, int b, uint8_t *wc_img, float *r64_out, cudaStream_t st);
int heads_tc_cls(pcp_model *m, const float *g3, int b, float *d_cls, float *d_cls_logits, cudaStream_t st);
int heads_tc_seg(pcp_model *m, const float *d_x, int b, int64_t n, const float *centroid, const float *scale, const float *d_R,
                 const float *d_R64, const float *g3, float *d_seg, float *d_seg_logits, cudaStream_t st);
void train_destroy(pcp_model *m);
int model_seg_head_fp32(pcp_model *m, const float *d_x, int b, int64_t n, const float *d_R, const float *d_R64, const float *g3,
                        float *d_seg, float *d_seg_logits, cudaStream_t st);
}
