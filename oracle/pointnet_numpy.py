"""ORACLE — TEST INFRASTRUCTURE ONLY.  NumPy fp32 restatement of the reference PointNet hot path.

This file is the checker, never the product: only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import it.  Nothing under
`pointcloudprocessing_b200/` imports `oracle/`.

PARITY: GRAPH PINNED, KERAS PRIMITIVES UNPINNED.  The reference (MAPieschl/PointCloudProcessing) ships no golden
vectors, no known-answer tests and no weights, and TensorFlow 2.20.0 / Keras 3.10.0
(`point_cloud_analysis/requirements.txt:46,113`) cannot be imported in this environment.  The reference's own model
code (pointnet/PointNet.py, unmodified) IS executed here on a minimal Keras stand-in by
tests/golden/make_network_golden.py, and tests/test_network_reference.py holds this restatement to its outputs
(inference, training-mode step with gradients, frozen-network forward): layer table, names, dataflow, T-Net algebra,
regulariser and freeze semantics are pinned.  What stays restated from the published Keras semantics on both sides:
BN eps 1e-3 / biased variance / `training and trainable`, Glorot-uniform, SCCE clip 1e-7, Adam eps, dropout scaling.

All citations are `path:line` under /root/reference/point_cloud_analysis/ (PCA/).
Everything is explicit float32.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32
BN_EPS = F32(1e-3)          # keras.layers.BatchNormalization default epsilon (PCA/pointnet/PointNet.py:528)
BN_MOMENTUM = F32(0.99)     # PCA/pointnet/PointNet.py:502,528

TNET_WIDTHS = (64, 128, 1024, 512, 256)   # PCA/pointnet/PointNet.py:384


# --------------------------------------------------------------------------------------------
# layer table (PCA/pointnet/PointNet.py:116-141, :406-416)
# --------------------------------------------------------------------------------------------
def conv_name(base: str) -> str:
    """ConvLayer appends '_convolution_layer' to its name (PointNet.py:529)."""
    return f"{base}_convolution_layer"


def dense_name(base: str) -> str:
    """DenseLayer appends '_dense_layer' to its name (PointNet.py:625)."""
    return f"{base}_dense_layer"


def tnet_layers(name: str, k: int):
    """(kind, layer name, cin, cout) for one TNet (PointNet.py:406-410)."""
    w = TNET_WIDTHS
    return [
        ("conv", conv_name(f"{name}_convolution_layer_1"), k, w[0]),
        ("conv", conv_name(f"{name}_convolution_layer_2"), w[0], w[1]),
        ("conv", conv_name(f"{name}_convolution_layer_3"), w[1], w[2]),
        ("dense", dense_name(f"{name}_dense_layer_1"), w[2], w[3]),
        ("dense", dense_name(f"{name}_dense_layer_2"), w[3], w[4]),
    ]


def layer_table(C: int, P: int, vanilla: bool = False):
    """Every weighted layer: (kind, name, cin, cout, has_bn).  Order = PointNet.py:116-141."""
    t = []
    if not vanilla:
        t += [(k, n, ci, co, True) for (k, n, ci, co) in tnet_layers("input_transform", 3)]
    t += [("conv", conv_name("s1_l1_64"), 3, 64, True), ("conv", conv_name("s1_l2_64"), 64, 64, True)]
    if not vanilla:
        t += [(k, n, ci, co, True) for (k, n, ci, co) in tnet_layers("feature_transform", 64)]
    t += [
        ("conv", conv_name("s2_l1_64"), 64, 64, True),
        ("conv", conv_name("s2_l2_128"), 64, 128, True),
        ("conv", conv_name("s2_l3_1024"), 128, 1024, True),
        ("dense", dense_name("s3_l1_512"), 1024, 512, True),
        ("dense", dense_name("s3_l2_256"), 512, 256, True),
        ("dense", dense_name("output"), 256, C, False),          # use_bias = not apply_bn (PointNet.py:630)
        ("conv", conv_name("seg_l1_512"), 1088, 512, True),
        ("conv", conv_name("seg_l2_256"), 512, 256, True),
        ("conv", conv_name("seg_l3_128"), 256, 128, True),
        ("conv", conv_name("seg_l4_128"), 128, 128, True),
        ("conv", conv_name("seg_l5_output"), 128, P, False),      # PointNet.py:141, use_bias (:540)
    ]
    return t


def glorot_uniform(rng, cin: int, cout: int) -> np.ndarray:
    """keras GlorotUniform: U(-l, l), l = sqrt(6 / (fan_in + fan_out)) (PointNet.py:532,627)."""
    lim = np.sqrt(6.0 / (cin + cout))
    return rng.uniform(-lim, lim, size=(cin, cout)).astype(F32)


def make_params(C: int, P: int, seed: int = 7, vanilla: bool = False, trained_like: bool = True):
    """Generate a full parameter dict keyed by the reference's layer names.

    trained_like=True draws non-trivial BN statistics (gamma U(0.5,1.5), beta N(0,0.1),
    mean N(0,0.1), var U(0.5,1.5)) so the folded inference affine is exercised;
    trained_like=False gives the Keras build-time state gamma=1, beta=0, mean=0, var=1 (PointNet.py:528).
    Conv kernels are stored (Cin, Cout) (Keras (1,1,Cin,Cout) squeezed), dense (Cin, Cout).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    p = {}
    for kind, name, cin, cout, has_bn in layer_table(C, P, vanilla):
        p[f"{name}/kernel"] = glorot_uniform(rng, cin, cout)
        if has_bn:
            if trained_like:
                p[f"{name}/bn/gamma"] = rng.uniform(0.5, 1.5, cout).astype(F32)
                p[f"{name}/bn/beta"] = (0.1 * rng.standard_normal(cout)).astype(F32)
                p[f"{name}/bn/moving_mean"] = (0.1 * rng.standard_normal(cout)).astype(F32)
                p[f"{name}/bn/moving_variance"] = rng.uniform(0.5, 1.5, cout).astype(F32)
            else:
                p[f"{name}/bn/gamma"] = np.ones(cout, F32)
                p[f"{name}/bn/beta"] = np.zeros(cout, F32)
                p[f"{name}/bn/moving_mean"] = np.zeros(cout, F32)
                p[f"{name}/bn/moving_variance"] = np.ones(cout, F32)
        else:
            p[f"{name}/bias"] = (0.05 * rng.standard_normal(cout)).astype(F32) if trained_like else np.zeros(cout, F32)
    if not vanilla:
        for tn, k in (("input_transform", 3), ("feature_transform", 64)):
            w = glorot_uniform(rng, 256, k * k)                      # PointNet.py:415
            if trained_like:
                w = (w * F32(0.25)).astype(F32)                      # keep predicted matrices near I, like a trained T-Net
            p[f"{tn}/w"] = w
            p[f"{tn}/b"] = np.eye(k, dtype=F32)                      # 'identity' initializer, PointNet.py:416
    return p


# --------------------------------------------------------------------------------------------
# ops
# --------------------------------------------------------------------------------------------
def normalize(pc: np.ndarray):
    """PointCloudNormalization.call (PointNet.py:691-706).  pc: (b, n, 3) float32."""
    pc = np.asarray(pc, F32)
    # :694 tf.reduce_mean — TF's reduction order is unknowable here; the oracle takes the correctly
    # rounded mean (float64 accumulate, round once) and tests allow fp32 reduction-order slack.
    centroid = np.mean(pc, axis=1, keepdims=True, dtype=np.float64).astype(F32)
    centered = (pc - centroid).astype(F32)                               # :695
    dist = np.sqrt(np.sum(np.square(centered), axis=-1, dtype=F32))      # :696
    max_dist = np.max(dist, axis=1, keepdims=True)[..., None]            # :697-698
    scale = np.maximum(max_dist, F32(1e-7))                              # :701
    return (centered / scale).astype(F32), (centroid, scale)             # :704-706


def bn_inference(x, p, name):
    """keras BatchNormalization, training=False: (x - mean) / sqrt(var + eps) * gamma + beta."""
    g, b = p[f"{name}/bn/gamma"], p[f"{name}/bn/beta"]
    m, v = p[f"{name}/bn/moving_mean"], p[f"{name}/bn/moving_variance"]
    inv = (g / np.sqrt(v + BN_EPS)).astype(F32)
    return (x * inv + (b - m * inv)).astype(F32)


def bn_training(x, p, name, stats_out=None):
    """keras BatchNormalization, training=True and trainable: batch mean / biased variance over
    every axis but the last; moving stats updated with momentum 0.99 (biased batch variance)."""
    red = tuple(range(x.ndim - 1))
    mean = np.mean(x, axis=red, dtype=np.float64)
    var = np.var(x, axis=red, dtype=np.float64)
    g, b = p[f"{name}/bn/gamma"], p[f"{name}/bn/beta"]
    y = (x - mean.astype(F32)) * (g / np.sqrt(var.astype(F32) + BN_EPS)).astype(F32) + b
    if stats_out is not None:
        stats_out[name] = (mean.astype(F32), var.astype(F32))
    return y.astype(F32)


def fold_bn(p, name):
    """Inference BN as y = x*s + t (tf2onnx folds it the same way)."""
    g, b = p[f"{name}/bn/gamma"], p[f"{name}/bn/beta"]
    m, v = p[f"{name}/bn/moving_mean"], p[f"{name}/bn/moving_variance"]
    s = (g / np.sqrt(v + BN_EPS)).astype(F32)
    return s, (b - m * s).astype(F32)


def _layer(x, p, name, training, bn_train_layers, activation="relu", stats_out=None):
    """ConvLayer.call / DenseLayer.call (PointNet.py:554-566, :642-654): matmul (+bias) -> BN -> act."""
    y = (x @ p[f"{name}/kernel"]).astype(F32)
    if f"{name}/bias" in p:
        y = (y + p[f"{name}/bias"]).astype(F32)
    if f"{name}/bn/gamma" in p:
        use_batch = training and (bn_train_layers is None or name in bn_train_layers)
        y = bn_training(y, p, name, stats_out) if use_batch else bn_inference(y, p, name)
    if activation == "relu":
        y = np.maximum(y, F32(0))
    return y


def softmax(z):
    z = z - np.max(z, axis=-1, keepdims=True)
    e = np.exp(z, dtype=F32)
    return (e / np.sum(e, axis=-1, keepdims=True, dtype=F32)).astype(F32)


def tnet_forward(x, p, name, k, training=False, bn_train_layers=None, stats_out=None):
    """TNet.call (PointNet.py:418-454).  x: (b, n, k).  Returns (b, k, k) and the pooled feature."""
    names = [n for (_, n, _, _) in tnet_layers(name, k)]
    h = x
    for nm in names[:3]:
        h = _layer(h, p, nm, training, bn_train_layers, stats_out=stats_out)   # :423-425
    g = np.max(h, axis=1)                                                       # :429
    h = g
    for nm in names[3:]:
        h = _layer(h, p, nm, training, bn_train_layers, stats_out=stats_out)   # :431-432
    m = (h @ p[f"{name}/w"]).astype(F32).reshape(-1, k, k)                      # :434-439
    m = (m + p[f"{name}/b"]).astype(F32)                                        # :442
    return m, g


def tnet_reg_loss(m):
    """1e-3 * tf.nn.l2_loss(I - X X^T), summed over the batch (PointNet.py:447-451)."""
    k = m.shape[-1]
    d = np.eye(k, dtype=F32) - m @ np.transpose(m, (0, 2, 1))
    return F32(1e-3) * F32(0.5) * np.sum(np.square(d), dtype=F32)


def forward(pc, p, vanilla=False, training=False, bn_train_layers=None, dropout_masks=None,
            dropout_rate=0.3, return_intermediates=False, stats_out=None):
    """PointNet.call (PointNet.py:197-292).

    Returns dict with cls (b,C) probabilities, seg (b,n,P) probabilities, R (b,3,3), plus the
    pre-softmax logits (cls_logits, seg_logits) used for tolerance checks.
    dropout_masks: optional (mask1 (b,512), mask2 (b,256)) of 0/1 keep flags for training mode;
    kept units are scaled by 1/(1-rate) (Keras Dropout).
    """
    pcn, _ = normalize(pc)                                                       # :202
    b, n, _ = pcn.shape
    inter = {}
    if not vanilla:
        R, g1 = tnet_forward(pcn, p, "input_transform", 3, training, bn_train_layers, stats_out)  # :205
        X = (pcn @ R).astype(F32)                                                # :207
        inter["g_input_transform"] = g1
    else:
        R = np.tile(np.eye(3, dtype=F32), (b, 1, 1))                             # :211
        X = pcn
    X = _layer(X, p, conv_name("s1_l1_64"), training, bn_train_layers, stats_out=stats_out)   # :217
    X = _layer(X, p, conv_name("s1_l2_64"), training, bn_train_layers, stats_out=stats_out)   # :220
    if not vanilla:
        R64, g2 = tnet_forward(X, p, "feature_transform", 64, training, bn_train_layers, stats_out)  # :227
        X64 = (X @ R64).astype(F32)                                              # :228
        inter["g_feature_transform"] = g2
        inter["R64"] = R64
    else:
        X64 = X                                                                  # :231
    X = _layer(X64, p, conv_name("s2_l1_64"), training, bn_train_layers, stats_out=stats_out)  # :236
    X = _layer(X, p, conv_name("s2_l2_128"), training, bn_train_layers, stats_out=stats_out)   # :239
    X = _layer(X, p, conv_name("s2_l3_1024"), training, bn_train_layers, stats_out=stats_out)  # :242
    gf = np.max(X, axis=1)                                                       # :248
    inter["global_features"] = gf

    h = _layer(gf, p, dense_name("s3_l1_512"), training, bn_train_layers, stats_out=stats_out)   # :252
    if training and dropout_masks is not None:
        h = (h * dropout_masks[0] * F32(1.0 / (1.0 - dropout_rate))).astype(F32)  # :255
    h = _layer(h, p, dense_name("s3_l2_256"), training, bn_train_layers, stats_out=stats_out)     # :257
    if training and dropout_masks is not None:
        h = (h * dropout_masks[1] * F32(1.0 / (1.0 - dropout_rate))).astype(F32)  # :260
    cls_logits = _layer(h, p, dense_name("output"), training, bn_train_layers, activation=None)   # :262
    cls = softmax(cls_logits)

    gt = np.broadcast_to(gf[:, None, :], (b, n, gf.shape[-1]))                   # :268-269
    s = np.concatenate([X64, gt], axis=-1)                                       # :270 order [X_64, global]
    for nm in ("seg_l1_512", "seg_l2_256", "seg_l3_128", "seg_l4_128"):          # :275-284
        s = _layer(s, p, conv_name(nm), training, bn_train_layers, stats_out=stats_out)
    seg_logits = _layer(s, p, conv_name("seg_l5_output"), training, bn_train_layers, activation=None)  # :287
    seg = softmax(seg_logits)
    out = {"cls": cls, "seg": seg, "R": R.astype(F32), "cls_logits": cls_logits, "seg_logits": seg_logits}
    if return_intermediates:
        out.update(inter)
    return out


# --------------------------------------------------------------------------------------------
# losses / metrics / optimizer (PCA/pointnet_train.py:310-351)
# --------------------------------------------------------------------------------------------
def scce(prob, labels):
    """keras SparseCategoricalCrossentropy(from_logits=False): clip to [1e-7, 1-1e-7], -log, mean."""
    pr = np.clip(np.take_along_axis(prob, labels[..., None].astype(np.int64), axis=-1)[..., 0], 1e-7, 1 - 1e-7)
    return F32(np.mean(-np.log(pr.astype(F32)), dtype=np.float64))


def mse(y, yhat):
    return F32(np.mean(np.square(y.astype(F32) - yhat.astype(F32)), dtype=np.float64))


def exp_decay_lr(lr0, step, decay_steps, decay_rate):
    """keras ExponentialDecay(staircase=False) (pointnet_train.py:310-315)."""
    return lr0 * decay_rate ** (step / decay_steps)


def adam_step(w, g, m, v, t, lr, b1=0.9, b2=0.999, eps=1e-7):
    """keras Adam (pointnet_train.py:317-319): t is the 1-based step count."""
    m[:] = b1 * m + (1 - b1) * g
    v[:] = b2 * v + (1 - b2) * g * g
    alpha = lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
    w[:] = w - alpha * m / (np.sqrt(v) + eps)
    return w, m, v


# --------------------------------------------------------------------------------------------
# preprocessing (PCA/pointcloud/PointCloudSet.py)
# --------------------------------------------------------------------------------------------
def jitter(x, noise, stdev3):
    """_parse_function jitter (PointCloudSet.py:335-337): x + noise * stdev (fp32 mul, then add)."""
    return (np.asarray(x, F32) + (np.asarray(noise, F32) * np.asarray(stdev3, F32)).astype(F32)).astype(F32)


def adjust_to_input_width(obs, labels, width, pad_indices=None):
    """_adjust_to_input_width (PointCloudSet.py:443-470): truncate to the first `width` points, or
    append obs[floor(U(0,N))] duplicates.  pad_indices (length width-N) replaces the global-RNG draw."""
    n = obs.shape[0]
    if n > width:
        return obs[:width], labels[:width]
    idx = np.asarray(pad_indices, np.int64)
    assert idx.shape[0] == width - n
    return np.concatenate([obs, obs[idx]], 0), np.concatenate([labels, labels[idx]], 0)


def get_dcm(roll_deg, pitch_deg, yaw_deg):
    """utils/mat_ops.py:6-39 — R = (roll . pitch . yaw)^T with the reference's sign conventions."""
    y, pt, r = np.deg2rad(yaw_deg), np.deg2rad(pitch_deg), np.deg2rad(roll_deg)
    Ry = np.array([[np.cos(y), np.sin(y), 0.0], [-np.sin(y), np.cos(y), 0.0], [0.0, 0.0, 1.0]])
    Rp = np.array([[np.cos(pt), 0.0, -np.sin(pt)], [0.0, 1.0, 0.0], [np.sin(pt), 0.0, np.cos(pt)]])
    Rr = np.array([[1.0, 0.0, 0.0], [0.0, np.cos(r), np.sin(r)], [0.0, -np.sin(r), np.cos(r)]])
    return (Rr @ (Rp @ (Ry @ np.eye(3)))).T


def rotate(x, R):
    """Rotation augmentation (not in the reference): y = x . R^T per cloud, fp32,
    evaluated as ((x0*r0 + x1*r1) + x2*r2) with separate roundings (no FMA)."""
    x = np.asarray(x, F32)
    R = np.asarray(R, F32)
    out = np.empty_like(x)
    for j in range(3):
        a = (x[..., 0] * R[:, None, j, 0]).astype(F32)
        bb = (x[..., 1] * R[:, None, j, 1]).astype(F32)
        c = (x[..., 2] * R[:, None, j, 2]).astype(F32)
        out[..., j] = ((a + bb).astype(F32) + c).astype(F32)
    return out


def parse_aftr_frame_lines(text: str, class_labels: dict, part_labels: dict):
    """ORACLE (timed CPU baseline of the ingest path): the per-line loop of PointCloudSet.add_from_aftr_output
    (PCA/pointcloud/PointCloudSet.py:161-198) restated for one frame's text: strip, split the "(x, y, z)" part on ',',
    float() each value, split the labels on ' ', np.isfinite per point.  Returns (obs (N,3) float64, class index, part (N,))."""
    obs, cl, pl = [], None, []
    for line in text.splitlines():
        line = line.strip()
        a, b = line.find('('), line.find(')')
        pos = [float(v) for v in line[a + 1:b].split(',')]
        labels = [l for l in line[b + 1:].split(" ") if len(l) > 1]
        if len(labels) == 2:
            if labels[0] not in class_labels:
                raise Exception(f"Class label {labels[0]} not known")
            if labels[1] not in part_labels:
                raise Exception(f"Part label {labels[1]} not known")
        else:
            raise Exception(f"Dataset must contain both a class label and part label. {labels} is not correct.")
        if np.isfinite(np.array(pos)).all():
            obs.append(np.array(pos))
            cl = class_labels.get(labels[0], -1)
            pl.append(part_labels.get(labels[1], -1))
    return np.array(obs), cl, np.array(pl)
