"""ORACLE — TEST INFRASTRUCTURE ONLY.  PyTorch-CPU twin of oracle/pointnet_numpy.py.

Same graph (PCA/pointnet/PointNet.py:197-292), same parameter dict, evaluated with torch CPU fp32
so that (a) autograd pins the backward pass / training step of PCA/pointnet_train.py:310-351 and
(b) the forward can be timed on all host cores as the CPU baseline (MKL/oneDNN), standing in for
the reference's TensorFlow CPU path, which cannot be installed here.
PARITY UNPINNED (see oracle/pointnet_numpy.py).  Never imported by the product package.
"""
from __future__ import annotations

import numpy as np
import torch

from . import pointnet_numpy as onp

BN_EPS = 1e-3


def to_torch(params: dict, requires_grad=False) -> dict:
    out = {}
    for k, v in params.items():
        t = torch.from_numpy(np.array(v, np.float32, copy=True))
        if requires_grad and "moving_" not in k:
            t.requires_grad_(True)
        out[k] = t
    return out


def normalize(pc):
    """PointNet.py:691-706 (centroid in float64 like the NumPy oracle)."""
    centroid = pc.double().mean(dim=1, keepdim=True).float()
    c = pc - centroid
    dist = torch.sqrt((c * c).sum(-1))
    scale = torch.clamp(dist.max(dim=1, keepdim=True).values[..., None], min=1e-7)
    return c / scale


_GEMM_BF16 = False


class bf16_gemms:
    """Context manager: model the product's precision="bf16" TRAINING arithmetic — the operands of every layer GEMM whose
    contraction length is a multiple of 64 are rounded to bf16 (straight-through for autograd), products and sums stay
    fp32.  The 3-channel input layers and the per-cloud pc@R / X@R64 products stay fp32, as on the device."""
    def __enter__(self):
        global _GEMM_BF16
        self._old, _GEMM_BF16 = _GEMM_BF16, True

    def __exit__(self, *a):
        global _GEMM_BF16
        _GEMM_BF16 = self._old


def _mm(x, w):
    if _GEMM_BF16 and x.shape[-1] % 64 == 0:
        x, w = x.bfloat16().float(), w.bfloat16().float()
    return x @ w


def _layer(x, p, name, training, bn_train_layers, act=True, new_stats=None):
    """PointNet.py:554-566 / :642-654."""
    y = _mm(x, p[f"{name}/kernel"])
    if f"{name}/bias" in p:
        y = y + p[f"{name}/bias"]
    if f"{name}/bn/gamma" in p:
        g, b = p[f"{name}/bn/gamma"], p[f"{name}/bn/beta"]
        if training and (bn_train_layers is None or name in bn_train_layers):
            red = tuple(range(y.dim() - 1))
            mean = y.mean(dim=red)
            var = y.var(dim=red, unbiased=False)
            if new_stats is not None:
                m0, v0 = p[f"{name}/bn/moving_mean"], p[f"{name}/bn/moving_variance"]
                new_stats[f"{name}/bn/moving_mean"] = (0.99 * m0 + 0.01 * mean.detach())
                new_stats[f"{name}/bn/moving_variance"] = (0.99 * v0 + 0.01 * var.detach())
        else:
            mean, var = p[f"{name}/bn/moving_mean"], p[f"{name}/bn/moving_variance"]
        y = (y - mean) * (g / torch.sqrt(var + BN_EPS)) + b
    return torch.relu(y) if act else y


def _tnet(x, p, name, k, training, bnl, new_stats):
    """PointNet.py:418-454."""
    names = [n for (_, n, _, _) in onp.tnet_layers(name, k)]
    h = x
    for nm in names[:3]:
        h = _layer(h, p, nm, training, bnl, new_stats=new_stats)
    h = torch.amax(h, dim=1)            # even gradient split on ties, like tf.reduce_max
    for nm in names[3:]:
        h = _layer(h, p, nm, training, bnl, new_stats=new_stats)
    m = _mm(h, p[f"{name}/w"]).reshape(-1, k, k) + p[f"{name}/b"]
    return m


def forward(pc, p, vanilla=False, training=False, bn_train_layers=None, dropout_masks=None,
            dropout_rate=0.3, new_stats=None, heads=("cls", "seg")):
    """PointNet.call (PointNet.py:197-292).  pc: torch (b,n,3)."""
    cn, dn = onp.conv_name, onp.dense_name
    bnl = bn_train_layers
    pcn = normalize(pc)
    b, n, _ = pcn.shape
    out = {}
    if not vanilla:
        R = _tnet(pcn, p, "input_transform", 3, training, bnl, new_stats)
        X = pcn @ R
    else:
        R = torch.eye(3).repeat(b, 1, 1)
        X = pcn
    X = _layer(X, p, cn("s1_l1_64"), training, bnl, new_stats=new_stats)
    X = _layer(X, p, cn("s1_l2_64"), training, bnl, new_stats=new_stats)
    if not vanilla:
        R64 = _tnet(X, p, "feature_transform", 64, training, bnl, new_stats)
        X64 = X @ R64
        out["R64"] = R64
    else:
        X64 = X
    X = _layer(X64, p, cn("s2_l1_64"), training, bnl, new_stats=new_stats)
    X = _layer(X, p, cn("s2_l2_128"), training, bnl, new_stats=new_stats)
    X = _layer(X, p, cn("s2_l3_1024"), training, bnl, new_stats=new_stats)
    gf = torch.amax(X, dim=1)
    out["R"] = R
    out["global_features"] = gf
    if "cls" in heads:
        h = _layer(gf, p, dn("s3_l1_512"), training, bnl, new_stats=new_stats)
        if training and dropout_masks is not None:
            h = h * dropout_masks[0] * (1.0 / (1.0 - dropout_rate))
        h = _layer(h, p, dn("s3_l2_256"), training, bnl, new_stats=new_stats)
        if training and dropout_masks is not None:
            h = h * dropout_masks[1] * (1.0 / (1.0 - dropout_rate))
        out["cls_logits"] = _layer(h, p, dn("output"), training, bnl, act=False)
        out["cls"] = torch.softmax(out["cls_logits"], -1)
    if "seg" in heads:
        s = torch.cat([X64, gf[:, None, :].expand(b, n, gf.shape[-1])], -1)
        for nm in ("seg_l1_512", "seg_l2_256", "seg_l3_128", "seg_l4_128"):
            s = _layer(s, p, cn(nm), training, bnl, new_stats=new_stats)
        out["seg_logits"] = _layer(s, p, cn("seg_l5_output"), training, bnl, act=False)
        out["seg"] = torch.softmax(out["seg_logits"], -1)
    return out


def scce(prob, labels):
    """keras SparseCategoricalCrossentropy(from_logits=False) (pointnet_train.py:337-338)."""
    pr = torch.gather(prob, -1, labels[..., None].long())[..., 0].clamp(1e-7, 1 - 1e-7)
    return (-torch.log(pr)).mean()


def total_loss(out, y_cls, y_seg, y_se3, w_cls, w_seg, w_rot, reg_in=False, reg_feat=False):
    """compile(loss=..., loss_weights=...) (pointnet_train.py:334-345) + TNet add_loss (PointNet.py:447-451)."""
    l_cls = scce(out["cls"], y_cls)
    l_seg = scce(out["seg"], y_seg)
    l_rot = ((out["R"] - y_se3) ** 2).mean()
    loss = w_cls * l_cls + w_seg * l_seg + w_rot * l_rot
    for flag, key in ((reg_in, "R"), (reg_feat, "R64")):
        if flag and key in out:
            m = out[key]
            d = torch.eye(m.shape[-1]) - m @ m.transpose(1, 2)
            loss = loss + 1e-3 * 0.5 * (d * d).sum()
    return loss, (l_cls, l_seg, l_rot)


def adam_step_(p, grads, m, v, t, lr, b1=0.9, b2=0.999, eps=1e-7):
    """keras Adam update (pointnet_train.py:317-319), in place on dict-of-tensors."""
    alpha = lr * np.sqrt(1 - b2 ** t) / (1 - b1 ** t)
    with torch.no_grad():
        for k, g in grads.items():
            m[k].mul_(b1).add_(g, alpha=1 - b1)
            v[k].mul_(b2).addcmul_(g, g, value=1 - b2)
            p[k].sub_(alpha * m[k] / (v[k].sqrt() + eps))
