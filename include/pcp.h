/* pcp.h — C-ABI of libpcp_b200.so: the B200 (sm_100a) implementation of the PointNet point-cloud
 * hot path of MAPieschl/PointCloudProcessing.
 *
 * The reference has NO FFI / plugin / op-registration layer (it is 100 % Python on TensorFlow,
 * SURVEY.md §0.1, §8b): the boundary this library sits behind is the Keras Model/Layer contract
 * of point_cloud_analysis/pointnet/PointNet.py and the PointCloudSet preprocessing contract of
 * point_cloud_analysis/pointcloud/PointCloudSet.py.  Each entry point below cites the reference
 * symbol (file:line under /root/reference/point_cloud_analysis/) whose arithmetic it replaces.
 * The Python mirror of that contract (pointcloudprocessing_b200/pointnet, /pointcloud) binds these
 * symbols with ctypes (pointcloudprocessing_b200/_ffi.py); INTEGRATION.md shows the stub a
 * reference maintainer would add.
 *
 * Conventions
 *   - every function returns 0 (PCP_OK) or a negative pcp_status; pcp_last_error() gives the
 *     thread-local message; no C++ exception crosses the boundary;
 *   - plain pointers and sizes only.  Pointers named d_* are DEVICE pointers (e.g.
 *     torch.Tensor.data_ptr()), h_* are HOST pointers; the caller owns every tensor buffer;
 *   - launches are asynchronous on the cudaStream_t passed as `stream` (void*, 0 = default stream);
 *     the *_host entry points copy H2D/D2H themselves and synchronise before returning;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     PCP_ERR_CUDA;
 *   - EXCLUSIVE USE OF THE DEVICE is assumed by the kernels whose CTAs wait for each other inside one launch
 *     (pcp_normalize_f32 on LiDAR-sized clouds, pcp_fps_f32, the fused chain kernel of pcp_model_forward in bf16 mode):
 *     they size their grids to be fully resident, so under MPS / co-tenancy with another process's kernels a
 *     waited-for CTA may not be scheduled; every such wait is bounded and traps after ~2 s instead of hanging;
 *   - test / development hooks (pcp_selftest_*, pcp_bench_*, ...) are declared in pcp_dev.h, not here.
 */
#ifndef PCP_H_
#define PCP_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum pcp_status {
    PCP_OK = 0,
    PCP_ERR_INVALID = -1,      /* bad argument */
    PCP_ERR_CUDA = -2,         /* CUDA runtime error (message in pcp_last_error) */
    PCP_ERR_NOMEM = -3,
    PCP_ERR_NOTFOUND = -4,     /* unknown tensor name */
    PCP_ERR_UNSUPPORTED = -5,  /* shape outside what the kernels cover */
    PCP_ERR_STATE = -6         /* call order (e.g. forward before prepare) */
} pcp_status;

typedef void *pcp_stream;      /* cudaStream_t */

const char *pcp_last_error(void);
int pcp_version(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
uint64_t pcp_launch_count(void);
int pcp_device_info(int *sm_count, int *cc_major, int *cc_minor, size_t *total_mem);

/* ---- memory / stream plumbing (for callers without torch) ---- */
int pcp_malloc(void **d_ptr, size_t bytes);
int pcp_free(void *d_ptr);
int pcp_malloc_host(void **h_ptr, size_t bytes);          /* pinned */
int pcp_free_host(void *h_ptr);
int pcp_memcpy_h2d(void *d_dst, const void *h_src, size_t bytes, pcp_stream stream);
int pcp_memcpy_d2h(void *h_dst, const void *d_src, size_t bytes, pcp_stream stream);
int pcp_memset(void *d_ptr, int value, size_t bytes, pcp_stream stream);
int pcp_stream_sync(pcp_stream stream);

/* ---- preprocessing ---- */

/* PointCloudNormalization.call (pointnet/PointNet.py:691-706): per cloud centroid, centre,
 * max distance, scale = max(maxdist, 1e-7), divide.  x,y: (b, n, 3) f32 contiguous.
 * d_centroid (b*3) and d_scale (b) may be NULL; d_y may be NULL when both statistics are
 * requested (statistics-only pass).  One HBM read + one write (24 B/point).  Clouds of n >= 16384 points
 * (16-byte aligned, n % 4 == 0) run as shared-memory-resident slices that meet through per-cloud counters in scratch
 * memory (64 + 48 b bytes); there is no upper limit on n on that path.  pcp_normalize_f32 takes the scratch from a
 * library-owned per-device ring of eight slots (a slot is re-used only after the stream that used it last has passed its
 * kernel; growing the ring calls cudaFree / cudaMalloc); pcp_normalize_ws_f32 takes it from the caller — no allocation,
 * no hidden state — sized by pcp_normalize_workspace_bytes (0: this shape needs none).
 * Smaller or unaligned clouds: one thread-block cluster per cloud (n <= 273 066), no scratch. */
int pcp_normalize_f32(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale,
                      pcp_stream stream);
int pcp_normalize_workspace_bytes(int b, int64_t n, size_t *bytes);
int pcp_normalize_ws_f32(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale, void *d_workspace,
                         size_t workspace_bytes, pcp_stream stream);

/* Allocation-free variant of pcp_normalize_f32 for the sliced kernel (n >= 16384, n % 4 == 0, 16-byte aligned x / y).
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic and the same kernel as
 * pcp_normalize_f32; the per-cloud scratch (48 B per cloud + a 64-byte ticket block) is a static device array of
 * 64 + 4096 * 48 = 196 672 bytes that every GPU which loads the library holds until exit, so b is limited to
 * pcp_normalize_noalloc_max_clouds() = 4096 clouds per call (larger b, smaller or unaligned clouds: PCP_ERR_UNSUPPORTED
 * with a message; nothing is launched).  The call performs no cudaMalloc / cudaFree and calls no allocating library
 * routine; it clears the part of the array it uses on `stream` before the launch.  EVERY call on a device shares that one
 * array: two calls that overlap on the GPU (different streams, or graphs replayed concurrently) race — serialise them. */
int pcp_normalize_noalloc_max_clouds(void);
int pcp_normalize_noalloc_f32(const float *d_x, int b, int64_t n, float *d_y, float *d_centroid, float *d_scale,
                              pcp_stream stream);

/* PointCloudSet._parse_function jitter (pointcloud/PointCloudSet.py:335-337):
 * y = x + noise * stdev_axis.  d_noise != NULL: explicit noise tensor (parity tests).
 * d_noise == NULL: noise = Philox4x32-10 + Box-Muller, one Philox block per point, counter =
 * global point index + offset (shard-invariant).  npoints = b*n. */
int pcp_jitter_f32(const float *d_x, int64_t npoints, const float h_stdev[3], uint64_t seed, uint64_t offset,
                   const float *d_noise, float *d_y, pcp_stream stream);
/* the N(0,1) tensor pcp_jitter_f32 uses when d_noise == NULL: (npoints, 3) f32 */
int pcp_philox_normal_f32(int64_t npoints, uint64_t seed, uint64_t offset, float *d_noise, pcp_stream stream);

/* Rotation augmentation (not in the reference; DCM convention utils/mat_ops.py:6-39):
 * y = x . R^T per cloud, d_R (b, 3, 3). */
int pcp_rotate_f32(const float *d_x, int b, int64_t n, const float *d_R, float *d_y, pcp_stream stream);

/* PointCloudSet._adjust_to_input_width (pointcloud/PointCloudSet.py:443-470) on the device:
 * cloud c has h_n_in[c] valid points at d_x + c*n_in_stride*3; output is exactly n_out points:
 * the first n_out (truncate, :455-456) or all points followed by duplicates of
 * x[floor(U(0,N))] (:459-467) drawn from Philox(seed, cloud, slot).  Labels (int32, may be NULL)
 * are gathered with the same indices.  d_n_in: device int32[b]. */
int pcp_resize_to_width_f32(const float *d_x, const int32_t *d_labels, int b, int64_t n_in_stride,
                            const int32_t *d_n_in, int64_t n_out, uint64_t seed, float *d_y,
                            int32_t *d_labels_out, pcp_stream stream);

/* ---- sampling (north_star operators; no reference implementation, contract = oracle/sampling.py) ---- */

/* Farthest-point sampling: x (b, n, 3) -> idx (b, s) int32; first index = start; fp32
 * ((dx*dx + dy*dy) + dz*dz) without FMA, lowest index wins ties. */
int pcp_fps_f32(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, pcp_stream stream);
/* The same samples (bit-identical) with caller-owned scratch: with a workspace the library first puts every cloud in Morton order
 * of a coarse grid and then skips, per iteration, the groups of points whose bounding sphere is farther from the new sample than
 * their largest running distance; only the running distances stay on chip, so a 100 000-point cloud needs two thread blocks
 * instead of four — 1.9x faster on LiDAR-sized batches (64 x 100 000 -> 1024: 2.7 ms against 5.1 ms).
 * pcp_fps_workspace_bytes reports 0 where the plain kernel is the better choice (few or small clouds); pcp_fps_ws_f32 then
 * ignores the workspace and runs the plain kernel.  Clouds the plain kernel cannot keep on chip (n > 204 800) and that are too
 * few to be pruned stream their running distances through b * n floats of scratch: from the workspace here, from a
 * library-owned per-device buffer (cudaMalloc on growth) in pcp_fps_f32. */
int pcp_fps_workspace_bytes(int b, int64_t n, int s, size_t *bytes);
int pcp_fps_ws_f32(const float *d_x, int b, int64_t n, int s, int start, int32_t *d_idx, void *d_workspace, size_t workspace_bytes,
                   pcp_stream stream);
/* out[c, j, :] = x[c, idx[c, j], :] ; labels likewise when non-NULL */
int pcp_gather_points_f32(const float *d_x, const int32_t *d_labels, int b, int64_t n, const int32_t *d_idx, int s,
                          float *d_out, int32_t *d_labels_out, pcp_stream stream);

/* Voxel-grid downsample: key = floor((p - origin)/voxel) per axis; per-voxel centroid (double
 * accumulate) and count, voxels ordered lexicographically by (kx,ky,kz).  h_origin == NULL ->
 * per-cloud minimum corner.  Outputs are padded to n rows per cloud; d_nvox (b) int32 gives the
 * number of valid rows.  Workspace from pcp_voxel_workspace_bytes. */
int pcp_voxel_workspace_bytes(int b, int64_t n, size_t *bytes);
int pcp_voxel_f32(const float *d_x, int b, int64_t n, const float *h_origin, float voxel, float *d_out_xyz,
                  int32_t *d_out_count, int32_t *d_nvox, void *d_workspace, size_t workspace_bytes,
                  pcp_stream stream);

/* The same operator with caller-owned, capacity-bounded outputs, fully asynchronous: d_out_xyz (b, capacity, 3),
 * d_out_count (b, capacity); rows [0, min(nvox, capacity)) of a cloud are written, the others are left untouched (no
 * zero padding).  Nothing is allocated, nothing is read back and the call does not synchronise (workspace from
 * pcp_voxel_workspace_bytes).  d_nvox[c] = number of voxels of cloud c (may exceed capacity: the caller sees the
 * truncation), -1 for a cloud the contract rejects (non-finite coordinate), -2 for EVERY cloud when one cloud's bounding
 * grid exceeds 32 n cells (the bitmap path does not cover the batch): hand the batch to pcp_voxel_f32. */
int pcp_voxel_compact_f32(const float *d_x, int b, int64_t n, const float *h_origin, float voxel, int64_t capacity,
                          float *d_out_xyz, int32_t *d_out_count, int32_t *d_nvox, void *d_workspace,
                          size_t workspace_bytes, pcp_stream stream);

/* ---- model: PointNet (pointnet/PointNet.py:84-292) ---- */

typedef struct pcp_model pcp_model;

#define PCP_PRECISION_FP32 0   /* CUDA-core fp32 FMA path: the 1e-3 contract */
#define PCP_PRECISION_BF16 1   /* tcgen05 kind::f16, bf16 operands, fp32 accumulate in TMEM: the 2e-2 contract */
#define PCP_PRECISION_FP32_TC 2 /* the 1e-3 contract ON the tensor cores: every layer product evaluated as three bf16 tcgen05 products
                                  hi(A) hi(W) + hi(A) lo(W) + lo(A) hi(W) (hi = bf16(x), lo = bf16(x - hi): ~16 mantissa bits — more than
                                  TF32's 10), fp32 accumulation in TMEM, fp32 activations between layers.  Inference only: training
                                  steps of such a model run the fp32 CUDA-core path. */

#define PCP_HEAD_CLS 1u        /* classification_output */
#define PCP_HEAD_SEG 2u        /* segmentation_output   */

typedef struct pcp_model_desc {
    int32_t num_classes;               /* classification_output_width, PointNet.py:86 */
    int32_t num_parts;                 /* segmentation_output_width,   PointNet.py:87 */
    int32_t vanilla;                   /* PointNet.py:91 */
    int32_t regularize_input_transform;    /* PointNet.py:92 */
    int32_t regularize_feature_transform;  /* PointNet.py:93 */
    float dropout_rate;                /* PointNet.py:88 */
    int32_t precision;                 /* PCP_PRECISION_* */
} pcp_model_desc;

int pcp_model_create(const pcp_model_desc *desc, pcp_model **out);
int pcp_model_destroy(pcp_model *m);

/* tensor table: names follow the reference's layer names ("<layer>/kernel", "<layer>/bias",
 * "<layer>/bn/gamma|beta|moving_mean|moving_variance", "input_transform/w", ...). */
int pcp_model_num_tensors(const pcp_model *m);
int pcp_model_tensor_info(const pcp_model *m, int index, char *name_buf, size_t name_buf_len, size_t *count,
                          int *trainable_group);
int pcp_model_set_tensor(pcp_model *m, const char *name, const float *h_data, size_t count);
int pcp_model_get_tensor(pcp_model *m, const char *name, float *h_data, size_t count);

/* Fold inference BN (y = x*s + t), build device weight images (fp32 and packed bf16 UMMA tiles).
 * Must be called after set_tensor and before forward; training steps re-run it lazily. */
int pcp_model_prepare(pcp_model *m);
/* (re)allocate the activation workspace for up to b clouds of n points (outside the hot path). */
int pcp_model_reserve(pcp_model *m, int b, int64_t n);

/* PointNet.call(pc, training=False) (PointNet.py:197-292).  d_x (b,n,3) f32.  Outputs (each may
 * be NULL): d_cls (b,C) probabilities, d_seg (b,n,P) probabilities, d_R (b,3,3), and the
 * pre-softmax logits.  heads: PCP_HEAD_* mask of the heads to evaluate (the reference always
 * evaluates both). */
int pcp_model_forward(pcp_model *m, const float *d_x, int b, int64_t n, uint32_t heads, float *d_cls, float *d_seg,
                      float *d_R, float *d_cls_logits, float *d_seg_logits, pcp_stream stream);
/* same, HOST buffers: H2D of x, forward, D2H of the requested outputs, stream sync. */
int pcp_model_predict_host(pcp_model *m, const float *h_x, int b, int64_t n, uint32_t heads, float *h_cls,
                           float *h_seg, float *h_R, pcp_stream stream);
/* intermediate pooled features of the last forward: which = 0 input T-Net, 1 feature T-Net,
 * 2 global features; (b,1024) f32 copied to the host (tests). */
int pcp_model_get_pooled(pcp_model *m, int which, int b, float *h_out);

/* Measurement hook (bench.py's roofline): when enabled, every fused chain-kernel launch is
 * bracketed by CUDA events on its own stream; pcp_model_chain_time drains them and returns the
 * summed device time and the launch count since the last call. */
int pcp_model_enable_timing(pcp_model *m, int flag);
int pcp_model_chain_time(pcp_model *m, float *ms_total, int *launches);

/* ---- pose post-processing, batched: the reference's host NumPy float64 helpers on the device;
 * fp32 in / out, fp64 arithmetic inside; 3x3 matrices row-major, angles in degrees as (roll, pitch, yaw) ---- */
/* get_dcm (PCA/utils/mat_ops.py:6-39): R = (roll . pitch . yaw)^T; d_rpy_deg (b x 3) -> d_R (b x 9) */
int pcp_dcm_from_rpy_f32(const float *d_rpy_deg, int b, float *d_R, pcp_stream stream);
/* get_roll_pitch_yaw_deg (mat_ops.py:30-36) */
int pcp_rpy_from_dcm_f32(const float *d_R, int b, float *d_rpy_deg, pcp_stream stream);
/* reorthogonalize (mat_ops.py:63-65): U V^T of the SVD; like the reference, a reflection stays a reflection */
int pcp_reorthogonalize_f32(const float *d_R, int b, float *d_out, pcp_stream stream);
/* solve_kabsch (PCT/utils/calibration.py:3-31): b pairs of (n x 3) point sets -> R (b x 9), t (b x 3) with
 * aligned = R . measured + t, and the rmse (b; may be null) */
int pcp_kabsch_f32(const float *d_truth, const float *d_measured, int b, int64_t n, float *d_R, float *d_t, float *d_rmse,
                   pcp_stream stream);

/* ---- training (pointnet_train.py:310-351) ---- */
/* Data-parallel BatchNorm.  When a callback is set and pcp_model_loss_backward is called with
 * global_batch > b, every batch-statistics reduction of the step (forward sums and backward sums of each thawed
 * BatchNormalization layer; the Gram matrix of the implicitly handled pooled layers) is summed over the ranks through it
 * before it is used, so the data-parallel step computes the single-device step on the concatenated batch (the reference
 * trains on one device, pointnet_train.py:221-228).  fn must all-reduce(sum) `count` elements of d_buf in place (dtype 0 =
 * float32, 1 = float64), ordered on `stream`, and return 0; it is called on the thread that runs the step.  Equal shard
 * sizes are assumed (rows of the concatenated batch = local rows * global_batch / b). */
typedef int (*pcp_allreduce_fn)(void *user, void *d_buf, size_t count, int dtype, pcp_stream stream);
int pcp_model_set_bn_allreduce(pcp_model *m, pcp_allreduce_fn fn, void *user);
#define PCP_GROUP_SHARED 0
#define PCP_GROUP_INPUT_TRANSFORM 1
#define PCP_GROUP_CLS_HEAD 2
#define PCP_GROUP_SEG_HEAD 3
/* PointNet.freeze_* / thaw_* (PointNet.py:294-342): frozen => no gradient and BN in inference mode */
int pcp_model_set_trainable(pcp_model *m, int group, int flag);

typedef struct pcp_train_config {
    float w_cls, w_seg, w_rot;     /* loss_weights, pointnet_train.py:340-344 */
    float lr0, decay_rate;         /* ExponentialDecay, pointnet_train.py:310-315 */
    int32_t decay_steps;
    float beta1, beta2, eps;       /* keras Adam defaults 0.9 / 0.999 / 1e-7 */
    uint64_t dropout_seed;         /* Philox key of the Dropout masks (vary it per step) */
    int64_t sample_offset;         /* global index of this shard's first cloud: masks are shard-count invariant */
    int32_t use_dropout_masks;     /* 1: d_mask1/d_mask2 given explicitly (parity tests) */
    int32_t skip_unweighted_heads; /* 1: skip the segmentation head when it is frozen and w_seg == 0 (its metrics read 0) */
} pcp_train_config;

/* One training step on the local shard: forward (training BN on thawed layers, dropout),
 * losses (SCCE on probabilities clip 1e-7, MSE on R), backward into the flat gradient buffer.
 * h_metrics (8 floats): loss, cls_loss, seg_loss, se3_loss, cls_acc, seg_acc, se3_rmse, reg_loss.
 * global_batch: the loss mean divides by this (data parallel: sum over ranks == global mean).
 * Once a configuration (shapes, loss weights, thawed groups, which labels are given) has run eagerly two or three times (until
 * the table of packed weight images is complete) the step is replayed as one CUDA graph on a stream of the library's own, forked from and joined to `stream` by events; the inputs are
 * copied into staging buffers first, so the caller's pointers may change from call to call.  Steps with a BatchNorm
 * all-reduce callback or explicit dropout masks stay eager; PCP_TRAIN_GRAPH=0 in the environment keeps every step eager. */
int pcp_model_loss_backward(pcp_model *m, const float *d_x, int b, int64_t n, const int32_t *d_y_cls,
                            const int32_t *d_y_seg, const float *d_y_se3, const pcp_train_config *cfg,
                            int global_batch, const float *d_mask1, const float *d_mask2, float *h_metrics,
                            pcp_stream stream);
/* keras Model.evaluate / the validation pass of fit (pointnet_train.py:221-228): inference-mode forward and the
 * same 8 metrics as pcp_model_loss_backward (both T-Net regularisers included when the model was built with them).
 * Label pointers may be NULL (that loss reads 0; NULL d_y_seg skips the segmentation head). */
int pcp_model_evaluate(pcp_model *m, const float *d_x, int b, int64_t n, const int32_t *d_y_cls, const int32_t *d_y_seg,
                       const float *d_y_se3, const pcp_train_config *cfg, float *h_metrics, pcp_stream stream);
/* flat fp32 gradient buffer (device) for the NCCL all-reduce done by the host side */
int pcp_model_grad_buffer(pcp_model *m, float **d_grads, size_t *count);
/* fused Adam + ExponentialDecay over the flat buffer; step is 0-based (lr uses step, bias
 * correction uses step+1), frozen tensors are skipped. */
int pcp_model_adam_step(pcp_model *m, const pcp_train_config *cfg, int64_t step, pcp_stream stream);
/* outputs of the last training-mode forward, copied to the host (tests / metrics): which = 0 cls probabilities
 * (b,C), 1 seg probabilities (b,n,P), 2 R (b,3,3), 3 R64 (b,64,64), 4 global features (b,1024) */
int pcp_model_get_train_output(pcp_model *m, int which, float *h_out, size_t count);
/* zero the Adam moments: a fresh optimizer is compiled for every training profile (pointnet_train.py:310-319) */
int pcp_model_reset_optimizer(pcp_model *m);
/* copy one gradient tensor to the host (tests) */
int pcp_model_get_grad(pcp_model *m, const char *name, float *h_data, size_t count);

/* ---- host-side data plumbing either side of the path (no GPU needed) ---- */

/* CRC32C (Castagnoli) and the TFRecord masked form ((crc >> 15 | crc << 17) + 0xa282ead8) used by
 * tf.io.TFRecordWriter (pointcloud/PointCloudSet.py:251-288). */
uint32_t pcp_crc32c(const void *h_data, size_t bytes);
uint32_t pcp_crc32c_masked(const void *h_data, size_t bytes);
/* Walk a TFRecord file image: payload offset / length of every record.  verify_crc: 0 none, 1 length CRCs,
 * 2 length + payload CRCs.  h_offsets / h_lengths may be NULL (count only). */
int pcp_tfrecord_scan(const uint8_t *h_buf, size_t bytes, int verify_crc, uint64_t *h_offsets, uint64_t *h_lengths,
                      size_t capacity, size_t *count);
/* Decode one tf.train.Example of the PointCloudSet schema (pointcloud/PointCloudSet.py:100-105, :306-323):
 * observations float[width*3], class_label int64, part_labels int64[width] (written as int32), se3 float[9]. */
int pcp_example_decode(const uint8_t *h_rec, size_t bytes, int64_t width, float *h_obs, int64_t *h_class_label,
                       int32_t *h_part_labels, float *h_se3);
/* `count` records decoded by `threads` host threads straight into one batch — the map(parse, AUTOTUNE) + batch() stage of
 * the reference's tf.data pipeline (pointcloud/PointCloudSet.py:347-363): h_obs (count, width, 3) f32, h_class_labels
 * (count) i32, h_part_labels (count, width) i32, h_se3 (count, 3, 3) f32; h_recs / h_lengths: payload pointer and length of
 * every record (e.g. into memory-mapped TFRecord files indexed by pcp_tfrecord_scan). */
int pcp_example_decode_batch(const uint8_t *const *h_recs, const uint64_t *h_lengths, int count, int64_t width, float *h_obs,
                             int32_t *h_class_labels, int32_t *h_part_labels, float *h_se3, int threads);
/* Parse one AftrBurner frame text ("(x, y, z) class part" per line) with the semantics of
 * PointCloudSet.add_from_aftr_output (pointcloud/PointCloudSet.py:161-198).  Label lists are '\n'-separated. */
int pcp_parse_aftr_frame(const char *h_text, size_t bytes, const char *class_labels_nl, const char *part_labels_nl,
                         double *h_xyz, int32_t *h_part, size_t capacity, size_t *count, int32_t *class_label,
                         size_t *h_nonfinite);

/* Provizio.to_aftr_frame (point_cloud_toolkit/utils/Provizio.py:126-145): the AftrBurner frame text, one line
 * "(x, y, z) label label ..." per point, every coordinate formatted like the reference's f-string (shortest round-trip
 * digits, Python's positional / scientific rules).  h_points: n x 3 float32 (is_f64 == 0) or float64; h_label_ids:
 * (n x labels_per_point) int32 indices into the '\n'-separated label_table_nl (labels_per_point == 0: no labels).  The text
 * goes to h_out (capacity bytes), its length to *len; PCP_ERR_NOMEM with the size needed in *len when capacity is too small. */
int pcp_format_aftr_frame(const void *h_points, int is_f64, size_t n, const int32_t *h_label_ids, int labels_per_point,
                          const char *label_table_nl, char *h_out, size_t capacity, size_t *len);

#ifdef __cplusplus
}
#endif
#endif /* PCP_H_ */
