"""GPU parity: PointNet forward (pointnet/PointNet.py:197-292) through the C-ABI vs the oracle.
Tolerances are BASELINE.json's: 1e-3 relative for the fp32 path, 2e-2 for bf16, identical arg-max."""
import numpy as np
import pytest
import torch

from oracle import pointnet_numpy as onp
from oracle import pointnet_torch as opt
from tests import synth

pytestmark = pytest.mark.gpu

TOL = {"fp32": 1e-3, "bf16": 2e-2, "fp32_tc": 1e-3}


def rel_err(a, ref):
    """max |a - ref| / max |ref| per tensor (near-zero entries make element-wise ratios meaningless)."""
    return float(np.abs(a.astype(np.float64) - ref.astype(np.float64)).max() / max(np.abs(ref).max(), 1e-30))


def argmax_agrees(logits, ref_logits, tol):
    """arg-max must match wherever the oracle's top-2 gap exceeds what the tolerance can move."""
    ref = ref_logits.reshape(-1, ref_logits.shape[-1]).astype(np.float64)
    got = logits.reshape(-1, logits.shape[-1])
    srt = np.sort(ref, axis=-1)
    gap = srt[:, -1] - srt[:, -2]
    decided = gap > 2 * tol * np.abs(ref).max()
    same = got.argmax(-1) == ref.argmax(-1)
    return bool(same[decided].all()), float(decided.mean()), float(same.mean())


def build(precision, vanilla=False, C=23, P=12, seed=7):
    from pointcloudprocessing_b200.pointnet import PointNet
    p = onp.make_params(C, P, seed=seed, vanilla=vanilla)
    m = PointNet(C, P, 0.3, 42, vanilla=vanilla, precision=precision)
    m.set_weights_dict(p)
    return m, p


def oracle_forward(x, p, vanilla):
    with torch.no_grad():
        o = opt.forward(torch.from_numpy(x), opt.to_torch(p), vanilla=vanilla)
    return {k: v.numpy() for k, v in o.items()}


@pytest.mark.parametrize("precision", ["fp32", "bf16", "fp32_tc"])
# b = 1 / 17 / 32: the fused small-batch dense tails (tail_fused.cu, b <= 32) with one row, a ragged row tile and a full one; b = 33: just past them
@pytest.mark.parametrize("vanilla,b,n", [(False, 4, 256), (True, 4, 256), (False, 32, 1024), (False, 3, 2048), (False, 2, 200), (False, 1, 300), (False, 17, 128),
                                         (False, 33, 128)])
def test_forward_matches_oracle(lib, precision, vanilla, b, n):
    m, p = build(precision, vanilla)
    x = synth.gaussian_clouds(1001, b, n)
    out = m.forward_with_logits(torch.from_numpy(x).cuda())
    got = {k: v.cpu().numpy() for k, v in out.items()}
    ref = oracle_forward(x, p, vanilla)
    tol = TOL[precision]
    for k in ("R", "cls_logits", "seg_logits", "cls", "seg"):
        assert rel_err(got[k], ref[k]) <= tol, (k, rel_err(got[k], ref[k]))
    ok, decided, same = argmax_agrees(got["cls_logits"], ref["cls_logits"], tol)
    assert ok, ("cls argmax", decided, same)
    ok, decided, same = argmax_agrees(got["seg_logits"], ref["seg_logits"], tol)
    assert ok and same > 0.9, ("seg argmax", decided, same)
    assert np.allclose(got["cls"].sum(-1), 1.0, atol=1e-5) and np.allclose(got["seg"].sum(-1), 1.0, atol=1e-5)


@pytest.mark.parametrize("precision", ["fp32", "bf16", "fp32_tc"])
def test_pooled_features_and_golden(lib, precision):
    from pathlib import Path
    g = np.load(Path(__file__).parent / "golden" / "pointnet_small.npz")
    m, p = build(precision, False, int(g["C"]), int(g["P"]), int(g["weight_seed"]))
    x = synth.gaussian_clouds(int(g["input_seed"]), int(g["b"]), int(g["n"]))
    cls, seg, R = m(torch.from_numpy(x).cuda())
    tol = TOL[precision]
    assert rel_err(cls.cpu().numpy(), g["cls"]) <= tol
    assert rel_err(seg.cpu().numpy(), g["seg"]) <= tol
    assert rel_err(R.cpu().numpy(), g["R"]) <= tol
    assert rel_err(m.pooled_features(2, int(g["b"])), g["global_features"]) <= tol


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_host_path_equals_device_path(lib, precision):
    m, p = build(precision)
    x = synth.gaussian_clouds(4, 5, 512)
    c1, s1, r1 = m(torch.from_numpy(x).cuda())
    c2, s2, r2 = m(x)                                  # NumPy in -> pcp_model_predict_host (H2D/D2H inside the library)
    assert np.array_equal(c1.cpu().numpy(), c2) and np.array_equal(s1.cpu().numpy(), s2) and np.array_equal(r1.cpu().numpy(), r2)
    cls_only = m(torch.from_numpy(x).cuda(), heads=("cls",))
    assert cls_only[1] is None and np.array_equal(cls_only[0].cpu().numpy(), c2)


def test_reference_point_cloud_shapes(lib):
    """kc-46.txt (490 pts) / f-15_model.txt (313 pts)-shaped inputs padded to the network width like
    PointCloudSet._adjust_to_input_width, then classified: exercises n not a multiple of any tile."""
    m, p = build("fp32")
    for npts in (490, 313):
        x = synth.gaussian_clouds(npts, 2, npts)
        got = m.forward_with_logits(torch.from_numpy(x).cuda())
        ref = oracle_forward(x, p, False)
        assert rel_err(got["cls_logits"].cpu().numpy(), ref["cls_logits"]) <= 1e-3
        assert rel_err(got["seg_logits"].cpu().numpy(), ref["seg_logits"]) <= 1e-3


def test_trainability_listing_matches_reference_log(lib):
    """models/kc46_lidar/log_*.log:191-206 — the 16-entry listing for the classification_pretrain profile."""
    m, _ = build("fp32")
    m.thaw_shared_network(); m.thaw_input_transform(); m.thaw_classification_head(); m.freeze_segmentation_head()
    t = m.get_layer_trainability()
    assert list(t) == ["input_normalization", "input_transform", "s1_l1_64_convolution_layer", "s1_l2_64_convolution_layer",
                       "feature_transform", "s2_l1_64_convolution_layer", "s2_l2_128_convolution_layer",
                       "s2_l3_1024_convolution_layer", "s3_l1_512_dense_layer", "s3_l2_256_dense_layer", "output_dense_layer",
                       "seg_l1_512_convolution_layer", "seg_l2_256_convolution_layer", "seg_l3_128_convolution_layer",
                       "seg_l4_128_convolution_layer", "seg_l5_output_convolution_layer"]
    assert [v for v in t.values()] == [False] + [True] * 10 + [False] * 5


@pytest.mark.parametrize("b,n", [(600, 1024), (40, 8192)])
def test_bf16_many_tiles_per_cta_matches_fp32_path(lib, b, n):
    """Persistent-CTA steady state (tens of tiles per CTA, ring/phase wrap-around): the tcgen05 path
    must agree with the fp32 CUDA-core path on the pooled global features and the class logits, and
    a few clouds are checked against the oracle directly."""
    mb, p = build("bf16")
    mf, _ = build("fp32")
    x = synth.gaussian_clouds(77, b, n)
    xd = torch.from_numpy(x).cuda()
    ob = mb.forward_with_logits(xd, heads=("cls",))
    of = mf.forward_with_logits(xd, heads=("cls",))
    assert rel_err(mb.pooled_features(2, b), mf.pooled_features(2, b)) <= 2e-2
    assert rel_err(ob["cls_logits"].cpu().numpy(), of["cls_logits"].cpu().numpy()) <= 2e-2
    assert rel_err(ob["R"].cpu().numpy(), of["R"].cpu().numpy()) <= 2e-2
    sel = [0, b // 2, b - 1]
    ref = oracle_forward(x[sel], p, False)
    assert rel_err(ob["cls_logits"].cpu().numpy()[sel], ref["cls_logits"]) <= 2e-2
