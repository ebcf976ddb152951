"""`.keras` archives of the reference's PointNet (pointnet_train.py:207-214 ModelCheckpoint(*.keras), :280-295 load_model, :257 stage
chaining) without Keras / h5py: a zip of `config.json`, `metadata.json` and `model.weights.h5`.

Layout of `model.weights.h5` as Keras 3 (`keras/src/saving/saving_lib.py`, 3.10) writes it for the reference's classes: every
saveable object gets a group `<path>/vars` holding its OWN variables as datasets "0", "1", ... in `trainable + non_trainable`
order, and the path of a child is its parent's path joined with the ATTRIBUTE NAME under which the parent holds it
(PCA/pointnet/PointNet.py:116-141 for PointNet, :406-416 for TNet, :528-541 ConvLayer.conv / .bn, :623-630 DenseLayer.dense / .bn):

    mlp_1_1/conv/vars/0                      Conv2D kernel (1, 1, Cin, Cout)      [/1 bias when there is no BatchNormalization]
    mlp_1_1/bn/vars/0..3                     gamma, beta, moving_mean, moving_variance
    mlp_cls_3/dense/vars/0, 1                Dense kernel (Cin, Cout), bias
    input_transform/vars/0, 1                TNet.w (256, K*K), TNet.b (K, K)
    input_transform/conv_layer_1/conv/vars/0 ...

UNPINNED: neither Keras nor an archive written by it is available here (and the reference's trained blobs are missing), so the
layout above is restated from the Keras sources, not checked against a file; tests cover our writer against our reader, the
name mapping against the reference's attribute names (parsed from its source) and a save -> load -> identical-outputs round trip.
"""
from __future__ import annotations

import datetime
import io
import json
import os
import tempfile
import zipfile

import numpy as np

from . import hdf5_min

# attribute name in the reference's PointNet -> (kind, base layer name): PCA/pointnet/PointNet.py:116-141
POINTNET_ATTRS = [
    ("mlp_1_1", "conv", "s1_l1_64"), ("mlp_1_2", "conv", "s1_l2_64"),
    ("mlp_2_1", "conv", "s2_l1_64"), ("mlp_2_2", "conv", "s2_l2_128"), ("mlp_2_3", "conv", "s2_l3_1024"),
    ("mlp_cls_1", "dense", "s3_l1_512"), ("mlp_cls_2", "dense", "s3_l2_256"), ("mlp_cls_3", "dense", "output"),
    ("mlp_seg_1", "conv", "seg_l1_512"), ("mlp_seg_2", "conv", "seg_l2_256"), ("mlp_seg_3", "conv", "seg_l3_128"),
    ("mlp_seg_4", "conv", "seg_l4_128"), ("mlp_seg_5", "conv", "seg_l5_output"),
]
TNET_ATTRS = [("conv_layer_1", "conv", "convolution_layer_1"), ("conv_layer_2", "conv", "convolution_layer_2"), ("conv_layer_3", "conv", "convolution_layer_3"),
              ("dense_layer_1", "dense", "dense_layer_1"), ("dense_layer_2", "dense", "dense_layer_2")]
# saveables without variables still get an (empty) vars group
EMPTY_ATTRS = ["normalize_input"]
BN_ORDER = ["gamma", "beta", "moving_mean", "moving_variance"]


def _layer_entries(vanilla: bool):
    """[(h5 path prefix, kind, our layer name)] for every weighted layer of the model"""
    out = [(attr, kind, f"{base}_{'convolution' if kind == 'conv' else 'dense'}_layer") for attr, kind, base in POINTNET_ATTRS]
    if not vanilla:
        for t in ("input_transform", "feature_transform"):
            for attr, kind, base in TNET_ATTRS:
                out.append((f"{t}/{attr}", kind, f"{t}_{base}_{'convolution' if kind == 'conv' else 'dense'}_layer"))
    return out


def weights_to_h5_tree(weights: dict, vanilla: bool):
    """our {tensor name: array} -> ({h5 dataset path: array}, [group paths])"""
    ds, groups = {}, ["vars"] + [f"{a}/vars" for a in EMPTY_ATTRS]
    for prefix, kind, layer in _layer_entries(vanilla):
        inner = "conv" if kind == "conv" else "dense"
        k = np.asarray(weights[f"{layer}/kernel"], np.float32)
        if k.ndim < 2:                               # flat (PointNet.get_weights_dict): Cout is the length of the layer's BN gamma / bias
            cout = np.asarray(weights[f"{layer}/bn/gamma"] if f"{layer}/bn/gamma" in weights else weights[f"{layer}/bias"]).size
            k = k.reshape(-1, cout)
        k = k.reshape((1, 1, k.shape[-2], k.shape[-1]) if kind == "conv" else (k.shape[-2], k.shape[-1]))
        groups += [f"{prefix}/vars", f"{prefix}/{inner}/vars"]
        ds[f"{prefix}/{inner}/vars/0"] = k
        if f"{layer}/bias" in weights:
            ds[f"{prefix}/{inner}/vars/1"] = np.asarray(weights[f"{layer}/bias"], np.float32)
        if f"{layer}/bn/gamma" in weights:
            groups.append(f"{prefix}/bn/vars")
            for i, nm in enumerate(BN_ORDER):
                ds[f"{prefix}/bn/vars/{i}"] = np.asarray(weights[f"{layer}/bn/{nm}"], np.float32)
    if not vanilla:
        for t in ("input_transform", "feature_transform"):
            kk = int(round(np.asarray(weights[f"{t}/b"]).size ** 0.5))
            ds[f"{t}/vars/0"] = np.asarray(weights[f"{t}/w"], np.float32).reshape(256, kk * kk)
            ds[f"{t}/vars/1"] = np.asarray(weights[f"{t}/b"], np.float32).reshape(kk, kk)
    return ds, groups


def h5_tree_to_weights(tree: dict, vanilla: bool) -> dict:
    """inverse of weights_to_h5_tree; conv kernels come back as (Cin, Cout)"""
    w = {}

    def need(path):
        if tree.get(path) is None:
            raise KeyError(f"model.weights.h5 has no dataset '{path}'")
        return tree[path]
    for prefix, kind, layer in _layer_entries(vanilla):
        inner = "conv" if kind == "conv" else "dense"
        k = need(f"{prefix}/{inner}/vars/0")
        w[f"{layer}/kernel"] = np.ascontiguousarray(k.reshape(k.shape[-2], k.shape[-1]), np.float32)
        if tree.get(f"{prefix}/bn/vars/0") is not None:
            for i, nm in enumerate(BN_ORDER):
                w[f"{layer}/bn/{nm}"] = np.ascontiguousarray(need(f"{prefix}/bn/vars/{i}"), np.float32)
        else:
            w[f"{layer}/bias"] = np.ascontiguousarray(need(f"{prefix}/{inner}/vars/1"), np.float32)
    if not vanilla:
        for t in ("input_transform", "feature_transform"):
            w[f"{t}/w"] = np.ascontiguousarray(need(f"{t}/vars/0"), np.float32)
            w[f"{t}/b"] = np.ascontiguousarray(need(f"{t}/vars/1"), np.float32)
    return w


def keras_config(config: dict, input_width) -> dict:
    """config.json as keras.saving.serialize_keras_object(model) lays it out for the registered class `Project>PointNet`
    (PointNet.py:83 register_keras_serializable(package="Project"); get_config :351-364 — which omits `vanilla`, quirk q2: ours keeps it)."""
    ref_keys = ["classification_output_width", "segmentation_output_width", "dropout_rate", "random_seed", "debugging",
                "regularize_input_transform", "regularize_feature_transform"]
    cfg = {"name": "point_net", "trainable": True, "dtype": {"module": "keras", "class_name": "DTypePolicy", "config": {"name": "float32"}, "registered_name": None}}
    cfg.update({k: config[k] for k in ref_keys})
    cfg["vanilla"] = bool(config.get("vanilla", False))
    return {"module": None, "class_name": "PointNet", "config": cfg, "registered_name": "Project>PointNet",
            "build_config": {"input_shape": [None, input_width, 3]}}


def save_keras(path: str, weights: dict, config: dict, input_width=None) -> None:
    vanilla = bool(config.get("vanilla", False))
    ds, groups = weights_to_h5_tree(weights, vanilla)
    with tempfile.TemporaryDirectory() as d:
        h5 = os.path.join(d, "model.weights.h5")
        hdf5_min.write_h5(h5, ds, groups)
        meta = {"keras_version": "3.10.0", "date_saved": datetime.datetime.now().strftime("%Y-%m-%d@%H:%M:%S")}
        with zipfile.ZipFile(path, "w", zipfile.ZIP_DEFLATED) as z:
            z.writestr("metadata.json", json.dumps(meta))
            z.writestr("config.json", json.dumps(keras_config(config, input_width)))
            z.write(h5, "model.weights.h5")


def load_keras(path: str):
    """-> (weights dict keyed by our tensor names, constructor config)"""
    with zipfile.ZipFile(path) as z:
        cfg = json.loads(z.read("config.json"))
        with tempfile.TemporaryDirectory() as d:
            z.extract("model.weights.h5", d)
            tree = hdf5_min.read_h5(os.path.join(d, "model.weights.h5"))
    c = cfg.get("config", {})
    keep = ["classification_output_width", "segmentation_output_width", "dropout_rate", "random_seed", "debugging", "vanilla",
            "regularize_input_transform", "regularize_feature_transform"]
    config = {k: c[k] for k in keep if k in c}
    # a reference-written archive has no `vanilla` key (q2): the weight file says which it is
    if "vanilla" not in config:
        config["vanilla"] = tree.get("input_transform/vars/0") is None
    return h5_tree_to_weights(tree, bool(config["vanilla"])), config
