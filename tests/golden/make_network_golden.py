"""Golden vectors for the network rows from the REFERENCE'S OWN MODEL CODE executing in this container:
point_cloud_analysis/pointnet/PointNet.py is imported unmodified on top of tests/golden/keras_standin.py (TensorFlow and
Keras are not installable here).  PointNet.__init__/build/call, TNet, ConvLayer, DenseLayer, PointCloudNormalization and
the freeze/thaw methods are the reference's; the Keras primitives underneath are the stand-in's (see its header for what
that leaves unpinned).  Every variable of the reference model is overwritten BY NAME from oracle.pointnet_numpy.make_params
(the script asserts that the reference's variables and the oracle's layer table correspond one to one, shapes included),
so the test side regenerates the weights from the seed instead of storing 17 MB of them.

Cases written to tests/golden/network_reference.npz:
  inf_*      inference forward, full model                       [cls, seg, R]
  van_*      inference forward, vanilla=True                     [cls, seg, R = I]
  trn_*/smo_* training=True forward (smo: BN beta + 6, every ReLU active), both T-Net regularisers on:  outputs, sum(model.losses), updated moving statistics,
             loss gradients (autograd through the reference graph) for small tensors + every variable's gradient norm
  frz_*      training=True forward with the shared network frozen (BN of frozen layers uses moving statistics)
  trainability_json   get_layer_trainability() after each step of a freeze/thaw sequence
The loss in trn_* is the compile() recipe of pointnet_train.py:334-345 written with torch in this script
(sparse categorical cross-entropy on probabilities with Keras' 1e-7 clip, MSE on se3, weights 1/1/0.5).

Run:  python tests/golden/make_network_golden.py        (needs /root/reference)"""
import importlib.util
import json
import sys
from pathlib import Path

import numpy as np
import torch

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parent.parent))
import keras_standin                                                     # noqa: E402
from oracle import pointnet_numpy as onp                                 # noqa: E402  (parameter generator + layer names only)

REF = Path("/root/reference")
C, P, B, N, SEED = 3, 4, 4, 128, 11
W_CLS, W_SEG, W_ROT = 1.0, 1.0, 0.5


def load_reference():
    keras_standin.install()
    path = next(REF.rglob("point_cloud_analysis/pointnet/PointNet.py"))
    spec = importlib.util.spec_from_file_location("ref_pointnet", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def weighted_layers(model):
    """Every ConvLayer / DenseLayer of the reference model, T-Net internals included, in construction order."""
    out = []
    for layer in model._custom_layers:
        if hasattr(layer, "conv_layer_1"):                                   # a TNet
            out += [layer.conv_layer_1, layer.conv_layer_2, layer.conv_layer_3, layer.dense_layer_1, layer.dense_layer_2]
        elif hasattr(layer, "conv") or hasattr(layer, "dense"):
            out.append(layer)
    return out


def variables(model):
    """name -> tensor for every variable, named as the oracle names them (<layer name>/kernel, .../bn/gamma, <tnet>/w ...)."""
    v = {}
    for L in weighted_layers(model):
        prim = L.conv if hasattr(L, "conv") else L.dense
        v[f"{L.name}/kernel"] = prim.kernel
        if prim.use_bias:
            v[f"{L.name}/bias"] = prim.bias
        if L.apply_bn:
            for k in ("gamma", "beta", "moving_mean", "moving_variance"):
                v[f"{L.name}/bn/{k}"] = getattr(L.bn, k)
    for layer in model._custom_layers:
        if hasattr(layer, "conv_layer_1"):
            v[f"{layer.name}/w"], v[f"{layer.name}/b"] = layer.w, layer.b
    return v


def smooth_params(p):
    """BN beta + 6: every ReLU active, so gradients are a smooth function of the weights (tight tolerances apply)."""
    return {k: (v + np.float32(6.0)).astype(np.float32) if k.endswith("/bn/beta") else v for k, v in p.items()}


def build_model(ref, vanilla, reg, smooth=False):
    m = ref.PointNet(C, P, dropout_rate=0.0, random_seed=3, vanilla=vanilla, regularize_input_transform=reg, regularize_feature_transform=reg)
    m._is_outermost = True
    with torch.no_grad():
        m(torch.zeros(B, N, 3))                                              # builds everything (auto-build, like keras)
    p = onp.make_params(C, P, seed=SEED, vanilla=vanilla, trained_like=True)
    if smooth:
        p = smooth_params(p)
    v = variables(m)
    assert sorted(v) == sorted(p), (sorted(set(v) ^ set(p)))               # the layer table, by name, both ways
    with torch.no_grad():
        for name, t in v.items():
            src = torch.from_numpy(p[name])
            assert tuple(t.shape)[-src.dim():] == tuple(src.shape) and t.numel() == src.numel(), (name, t.shape, src.shape)
            t.copy_(src.reshape(t.shape))
    order = [(L.name, "conv" if hasattr(L, "conv") else "dense") for L in weighted_layers(m)]
    assert order == [(n, k) for (k, n, _, _, _) in onp.layer_table(C, P, vanilla)], "construction order differs from the oracle's table"
    return m, v


def scce(prob, labels):
    pr = torch.gather(prob, -1, labels[..., None])[..., 0].clamp(1e-7, 1 - 1e-7)
    return (-torch.log(pr)).mean()


def main():
    ref = load_reference()
    rng = np.random.default_rng(SEED)
    pc = (rng.normal(0, 1, (B, N, 3)) * rng.uniform(0.5, 20, (B, 1, 1)) + rng.normal(0, 30, (B, 1, 3))).astype(np.float32)
    y_cls = rng.integers(0, C, B)
    y_seg = rng.integers(0, P, (B, N))
    y_se3 = np.stack([onp.get_dcm(*rng.uniform(-30, 30, 3)) for _ in range(B)]).astype(np.float32)
    out = {"pc": pc, "y_cls": y_cls, "y_seg": y_seg, "y_se3": y_se3, "meta": np.array(json.dumps(
        {"C": C, "P": P, "seed": SEED, "w_cls": W_CLS, "w_seg": W_SEG, "w_rot": W_ROT, "dropout_rate": 0.0}))}
    x = torch.from_numpy(pc)

    # inference, full and vanilla
    for tag, vanilla in (("inf", False), ("van", True)):
        m, _ = build_model(ref, vanilla, reg=False)
        with torch.no_grad():
            cls, seg, R = m(x, training=False)
        out[f"{tag}_cls"], out[f"{tag}_seg"], out[f"{tag}_R"] = cls.numpy(), seg.numpy(), R.numpy()
        if not vanilla:
            out["inf_R64"] = m.feature_transform._last_predicted.numpy()

    # training step forward + backward through the reference graph: ordinary weights (trn) and all-ReLUs-active weights (smo)
    for tag, smooth in (("trn", False), ("smo", True)):
        m, v = build_model(ref, False, reg=True, smooth=smooth)
        cls, seg, R = m(x, training=True)
        reg = sum(m.losses)
        l_cls, l_seg = scce(cls, torch.from_numpy(y_cls)), scce(seg, torch.from_numpy(y_seg))
        l_rot = ((R - torch.from_numpy(y_se3)) ** 2).mean()
        loss = W_CLS * l_cls + W_SEG * l_seg + W_ROT * l_rot + reg
        names = [n for n, t in v.items() if t.requires_grad]
        grads = dict(zip(names, torch.autograd.grad(loss, [v[n] for n in names])))
        out[tag + "_cls"], out[tag + "_seg"], out[tag + "_R"] = cls.detach().numpy(), seg.detach().numpy(), R.detach().numpy()
        out[tag + "_losses"] = np.array([float(t.detach()) for t in (loss, l_cls, l_seg, l_rot, reg)], np.float64)
        out[tag + "_grad_names"] = np.array(names)
        out[tag + "_grad_norms"] = np.array([float(grads[n].norm()) for n in names], np.float64)
        for n in ("output_dense_layer/kernel", "output_dense_layer/bias", "seg_l5_output_convolution_layer/kernel", "input_transform/b",
                  "input_transform/w", "s1_l1_64_convolution_layer/kernel", "s1_l1_64_convolution_layer/bn/gamma",
                  "feature_transform_dense_layer_2_dense_layer/bn/beta", "s2_l3_1024_convolution_layer/bn/gamma"):
            out[tag + "_grad:" + n] = grads[n].reshape(onp.make_params(C, P, seed=SEED)[n].shape).numpy()
        for n in ("s1_l1_64_convolution_layer", "s2_l3_1024_convolution_layer", "input_transform_convolution_layer_3_convolution_layer",
                  "feature_transform_dense_layer_1_dense_layer", "s3_l2_256_dense_layer", "seg_l4_128_convolution_layer"):
            out[f"{tag}_stat:{n}/bn/moving_mean"] = v[f"{n}/bn/moving_mean"].detach().numpy().copy()
            out[f"{tag}_stat:{n}/bn/moving_variance"] = v[f"{n}/bn/moving_variance"].detach().numpy().copy()

    # frozen shared network, training=True: frozen BN layers run on their moving statistics
    m, v = build_model(ref, False, reg=False)
    m.freeze_shared_network()
    with torch.no_grad():
        cls, seg, R = m(x, training=True)
    out["frz_cls"], out["frz_seg"], out["frz_R"] = cls.numpy(), seg.numpy(), R.numpy()
    out["frz_stat:s1_l1_64_convolution_layer/bn/moving_mean"] = v["s1_l1_64_convolution_layer/bn/moving_mean"].numpy().copy()     # unchanged
    out["frz_stat:seg_l1_512_convolution_layer/bn/moving_mean"] = v["seg_l1_512_convolution_layer/bn/moving_mean"].numpy().copy()  # updated

    # freeze / thaw bookkeeping (pure reference logic)
    m, _ = build_model(ref, False, reg=False)
    seq = ["freeze_shared_network", "freeze_segmentation_head", "thaw_input_transform", "freeze_classification_head", "thaw_shared_network",
           "thaw_segmentation_head", "freeze_input_transform", "thaw_classification_head"]
    trace = [["initial", m.get_layer_trainability()]]
    for call in seq:
        getattr(m, call)()
        trace.append([call, m.get_layer_trainability()])
    out["trainability_json"] = np.array(json.dumps(trace))

    dst = HERE / "network_reference.npz"
    np.savez_compressed(dst, **out)
    print("wrote", dst, dst.stat().st_size, "bytes")
    print("losses", out["trn_losses"], "cls[0]", out["inf_cls"][0], "R[0]", out["inf_R"][0].ravel())


if __name__ == "__main__":
    main()
