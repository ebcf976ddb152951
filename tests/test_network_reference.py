"""CPU: the oracle against OUTPUTS OF THE REFERENCE'S OWN MODEL CODE — tests/golden/network_reference.npz was produced by
executing the unmodified point_cloud_analysis/pointnet/PointNet.py (PointNet / TNet / ConvLayer / DenseLayer /
PointCloudNormalization: layer table, dataflow, T-Net algebra, regulariser, freeze logic) in this container on a minimal
Keras stand-in (tests/golden/keras_standin.py; generator tests/golden/make_network_golden.py).  This pins the oracle's
graph wiring, naming, ordering and freeze semantics to the reference source; the Keras primitives (1x1 conv = matmul, BN
formulas, eps, momentum) are restated from the Keras documentation on both sides and stay unpinned."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import pointnet_numpy as onp
from oracle import pointnet_torch as opt

G = np.load(Path(__file__).parent / "golden" / "network_reference.npz")
META = json.loads(str(G["meta"]))
C, P, SEED = META["C"], META["P"], META["seed"]

HEADS = [onp.dense_name(s) for s in ("s3_l1_512", "s3_l2_256", "output")] + \
        [onp.conv_name(s) for s in ("seg_l1_512", "seg_l2_256", "seg_l3_128", "seg_l4_128", "seg_l5_output")]


def close(a, ref, tol):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    return np.abs(a - ref).max() <= tol * max(1.0, np.abs(ref).max())


def test_inference_matches_reference_graph():
    for tag, vanilla in (("inf", False), ("van", True)):
        p = onp.make_params(C, P, seed=SEED, vanilla=vanilla)
        o = onp.forward(G["pc"], p, vanilla=vanilla, return_intermediates=True)
        assert close(o["cls"], G[f"{tag}_cls"], 2e-5) and close(o["seg"], G[f"{tag}_seg"], 2e-5) and close(o["R"], G[f"{tag}_R"], 2e-5), tag
        assert np.array_equal(o["cls"].argmax(-1), G[f"{tag}_cls"].argmax(-1))
        assert (o["seg"].argmax(-1) == G[f"{tag}_seg"].argmax(-1)).mean() > 0.999
        t = opt.forward(torch.from_numpy(G["pc"]), opt.to_torch(p), vanilla=vanilla)
        assert close(t["cls"].numpy(), G[f"{tag}_cls"], 2e-5) and close(t["seg"].numpy(), G[f"{tag}_seg"], 2e-5)
        if not vanilla:
            assert close(o["R64"], G["inf_R64"], 2e-5) and close(t["R64"].numpy(), G["inf_R64"], 2e-5)
    assert np.array_equal(G["van_R"], np.tile(np.eye(3, dtype=np.float32), (G["pc"].shape[0], 1, 1)))      # PointNet.py:211


def smooth_params(p):
    return {k: (v + np.float32(6.0)).astype(np.float32) if k.endswith("/bn/beta") else v for k, v in p.items()}


@pytest.mark.parametrize("tag", ["trn", "smo"])
def test_training_step_matches_reference_graph(tag):
    """trn: ordinary weights — the gradient of a ReLU network is discontinuous in its pre-activations, so two fp32
    evaluations that differ in summation order disagree on a few ReLU masks (tests/test_gpu_train.py header): 3e-2 / 6e-2.
    smo: BN beta + 6, every ReLU active — the wiring is checked to 1e-3."""
    smooth = tag == "smo"
    p = onp.make_params(C, P, seed=SEED)
    if smooth:
        p = smooth_params(p)
    tol_norm, tol_elem = (1e-3, 1e-3) if smooth else (3e-2, 6e-2)
    tp = opt.to_torch(p, requires_grad=True)
    stats = {}
    out = opt.forward(torch.from_numpy(G["pc"]), tp, training=True, dropout_masks=None, new_stats=stats)
    loss, parts = opt.total_loss(out, torch.from_numpy(G["y_cls"]), torch.from_numpy(G["y_seg"]), torch.from_numpy(G["y_se3"]),
                                 META["w_cls"], META["w_seg"], META["w_rot"], reg_in=True, reg_feat=True)
    want = G[tag + "_losses"]                                          # total, cls, seg, rot, sum(model.losses)
    total = float(loss.detach())
    assert abs(total - want[0]) <= 2e-5 * want[0]
    for got, ref in zip(parts, want[1:4]):
        assert abs(float(got.detach()) - ref) <= 2e-5 * max(1.0, ref)
    reg = total - sum(w * float(q.detach()) for w, q in zip((META["w_cls"], META["w_seg"], META["w_rot"]), parts))
    assert abs(reg - want[4]) <= 1e-4 * max(want[4], 1e-2 * want[0])    # 1e-3 * l2_loss(I - X X^T), both T-Nets (PointNet.py:447-451)
    # batch statistics over 4 rows in the dense BN layers amplify fp32 summation-order differences: 2e-4 here, 2e-5 in inference
    for k in ("cls", "seg", "R"):
        assert close(out[k].detach().numpy(), G[f"{tag}_{k}"], 2e-4), k
    loss.backward()
    names = [str(n) for n in G[tag + "_grad_names"]]
    assert sorted(names) == sorted(k for k in p if "moving_" not in k)  # what the reference graph holds as trainable
    gmax = G[tag + "_grad_norms"].max()
    for n, ref in zip(names, G[tag + "_grad_norms"]):
        g = tp[n].grad
        got = 0.0 if g is None else float(g.norm())
        # tensors whose true gradient is zero (a BN shift ahead of another batch-statistics BN) hold round-off only: absolute floor
        assert abs(got - ref) <= tol_norm * ref + 1e-6 * gmax, (n, got, ref)
    for key in G.files:
        if key.startswith(tag + "_grad:"):
            n = key.split(":", 1)[1]
            assert np.abs(tp[n].grad.numpy() - G[key]).max() <= tol_elem * np.abs(G[key]).max() + 1e-6 * gmax, n
        if key.startswith(tag + "_stat:"):
            n = key.split(":", 1)[1]
            assert close(stats[n].numpy(), G[key], 1e-5), n


def test_frozen_shared_network_matches_reference_graph():
    """freeze_shared_network (PointNet.py:306-314) freezes the input transform too; with training=True only the heads'
    BN layers use batch statistics and update their moving averages."""
    p = onp.make_params(C, P, seed=SEED)
    stats = {}
    out = opt.forward(torch.from_numpy(G["pc"]), opt.to_torch(p), training=True, bn_train_layers=set(HEADS), new_stats=stats)
    assert close(out["cls"].numpy(), G["frz_cls"], 2e-5) and close(out["seg"].numpy(), G["frz_seg"], 2e-5) and close(out["R"].numpy(), G["frz_R"], 2e-5)
    k = "s1_l1_64_convolution_layer/bn/moving_mean"
    assert k not in stats and np.array_equal(G["frz_stat:" + k], p[k])
    k = "seg_l1_512_convolution_layer/bn/moving_mean"
    assert close(stats[k].numpy(), G["frz_stat:" + k], 1e-5) and not np.allclose(G["frz_stat:" + k], p[k])
