"""CPU: pin the oracle.  The reference has no golden vectors / known-answer tests for this path
("parity unpinned"), so the pins are (a) the structural evidence the reference's
own training logs hold, (b) published known-answer vectors of the primitives we restate (Philox),
(c) hand-derived answers, (d) agreement of three independent restatements (NumPy, PyTorch, C)."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import pointnet_numpy as onp
from oracle import pointnet_torch as opt
from oracle import sampling as osm
from tests import synth


# ---- (a) structural pins from the reference's artefacts ----
def test_parameter_counts_match_reference_model():
    # 4 224 684 parameters, 4 210 476 trainable (BN moving stats = 14 208) for C=23, P=12
    p = onp.make_params(23, 12)
    assert sum(v.size for v in p.values()) == 4_224_684
    assert sum(v.size for k, v in p.items() if "moving_" not in k) == 4_210_476


def test_layer_multiset_matches_tf2onnx_dump():
    # models/kc46_lidar/log_*.log:2215 — Conv2D x16, MatMul x7 (dense), 2 TNet `w` matmuls
    t = onp.layer_table(23, 12, vanilla=False)
    assert sum(1 for l in t if l[0] == "conv") == 16
    assert sum(1 for l in t if l[0] == "dense") == 7
    assert [l for l in t if not l[4]] == [("dense", "output_dense_layer", 256, 23, False),
                                          ("conv", "seg_l5_output_convolution_layer", 128, 12, False)]


def test_mac_counts():
    # trunk 434 569 MAC/point (incl. pc@R 9 and X@R64 4096), per-cloud small-M 3 022 848
    t = {l[1]: l for l in onp.layer_table(23, 12)}
    per_point = sum(ci * co for (k, n, ci, co, bn) in t.values() if k == "conv" and not n.startswith("seg")) + 9 + 4096
    assert per_point == 434_569
    small = sum(ci * co for (k, n, ci, co, bn) in t.values() if k == "dense") + 256 * 9 + 256 * 4096
    assert small == 3_022_848


# ---- (b) published known-answer vectors ----
def test_philox4x32_10_known_answers():
    # Random123 kat_vectors: philox4x32 10 rounds
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
        ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
        ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
    ]
    for ctr, key, exp in kat:
        out = osm.philox4x32_10(np.array([ctr], np.uint32), np.array([key], np.uint32))[0]
        assert tuple(int(v) for v in out) == exp


# ---- (c) hand-derived answers ----
def test_fps_cube_corners():
    # unit cube: from corner 0 the farthest is the opposite corner (d^2 = 3), then the d^2 = 2 corners
    # by lowest index, ...
    cube = np.array([[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)], np.float32)
    idx = osm.fps(cube, 8)
    assert idx[0] == 0 and idx[1] == 7
    assert sorted(idx.tolist()) == list(range(8))
    # after {0,7} every remaining corner has min-distance 1 -> lowest index first
    assert idx[2] == 1


def test_fps_collinear_ties_lowest_index():
    pts = np.array([[0, 0, 0], [2, 0, 0], [-2, 0, 0], [1, 0, 0], [-1, 0, 0]], np.float32)
    # from 0: points 1 and 2 tie at d^2 = 4 -> index 1; then 2 (d^2 4); then 3,4 tie at 1 -> 3; then 4
    assert osm.fps(pts, 5).tolist() == [0, 1, 2, 3, 4]


def test_fps_identical_points_returns_zero_after_start():
    pts = np.ones((16, 3), np.float32)
    assert osm.fps(pts, 4, start=5).tolist() == [5, 0, 0, 0]


def test_voxel_lattice():
    g = np.array([[x, y, z] for x in range(3) for y in range(2) for z in range(2)], np.float32)
    pts = np.concatenate([g + 0.25, g + 0.75]).astype(np.float32)       # two points per unit voxel
    cent, cnt, keys = osm.voxel_downsample(pts, 1.0, origin=(0, 0, 0))
    assert cent.shape == (12, 3) and (cnt == 2).all()
    assert np.array_equal(keys, g.astype(np.int32))                        # lexicographic (kx,ky,kz)
    assert np.allclose(cent, g + 0.5)


def test_normalize_degenerate_clouds():
    one = np.array([[[3.0, -2.0, 5.0]]], np.float32)                       # single point: scale = 1e-7 path (PointNet.py:701)
    y, (c, s) = onp.normalize(one)
    assert np.array_equal(y, np.zeros_like(one)) and s.item() == np.float32(1e-7)
    same = np.tile(one, (1, 7, 1))
    y, (c, s) = onp.normalize(same)
    assert np.array_equal(y, np.zeros_like(same))
    x = synth.gaussian_clouds(1, 3, 500)
    y, _ = onp.normalize(x)
    assert np.allclose(np.linalg.norm(y, axis=-1).max(axis=1), 1.0, atol=1e-6)
    assert np.abs(y.mean(axis=1)).max() < 1e-5


def test_adjust_to_input_width_truncates_and_pads():
    obs = np.arange(30, dtype=np.float32).reshape(10, 3)
    lab = np.arange(10)
    o, l = onp.adjust_to_input_width(obs, lab, 4)
    assert np.array_equal(o, obs[:4]) and np.array_equal(l, lab[:4])       # PointCloudSet.py:455-456
    o, l = onp.adjust_to_input_width(obs, lab, 13, pad_indices=[9, 0, 3])
    assert o.shape == (13, 3) and np.array_equal(o[10:], obs[[9, 0, 3]]) and np.array_equal(l[10:], [9, 0, 3])


def test_get_dcm_is_rotation_and_matches_convention():
    R = onp.get_dcm(10.0, -20.0, 30.0)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and np.isclose(np.linalg.det(R), 1.0)
    assert np.allclose(onp.get_dcm(0, 0, 90), np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1]]), atol=1e-12)  # mat_ops.py:6-12,38-39


# ---- (d) independent restatements agree ----
@pytest.mark.parametrize("vanilla", [False, True])
def test_numpy_and_torch_oracles_agree(vanilla):
    p = onp.make_params(23, 12, seed=7, vanilla=vanilla)
    x = synth.gaussian_clouds(1001, 3, 200)
    a = onp.forward(x, p, vanilla=vanilla)
    b = opt.forward(torch.from_numpy(x), opt.to_torch(p), vanilla=vanilla)
    for k in ("cls", "seg", "R", "cls_logits", "seg_logits"):
        ref = a[k]
        assert np.abs(ref - b[k].numpy()).max() <= 2e-5 * max(1.0, np.abs(ref).max()), k


def test_training_mode_oracles_agree():
    p = onp.make_params(5, 4, seed=3)
    x = synth.gaussian_clouds(5, 4, 64)
    rng = np.random.default_rng(0)
    masks = ((rng.random((4, 512)) > 0.3).astype(np.float32), (rng.random((4, 256)) > 0.3).astype(np.float32))
    a = onp.forward(x, p, training=True, dropout_masks=masks)
    b = opt.forward(torch.from_numpy(x), opt.to_torch(p), training=True, dropout_masks=tuple(torch.from_numpy(m) for m in masks))
    for k in ("cls", "seg", "R"):
        assert np.abs(a[k] - b[k].numpy()).max() <= 5e-5, k


def test_c_oracle_matches_numpy_oracle(oracle_c):
    x = synth.lidar_scans(2002, 2, 3000)
    s = 64
    out = np.empty((2, s), np.int32)
    oracle_c.oracle_fps_batch_f32(x.ctypes.data_as(ctypes.c_void_p), 2, ctypes.c_int64(3000), s, 0, out.ctypes.data_as(ctypes.c_void_p))
    assert np.array_equal(out, osm.fps_batch(x, s))
    y = np.empty_like(x)
    oracle_c.oracle_normalize_f32(x.ctypes.data_as(ctypes.c_void_p), 2, ctypes.c_int64(3000), y.ctypes.data_as(ctypes.c_void_p), None, None)
    assert np.array_equal(y, onp.normalize(x)[0])
    oracle_c.oracle_voxel_f32.restype = ctypes.c_int64
    cx = np.empty((3000, 3), np.float32)
    cc = np.empty(3000, np.int32)
    v = oracle_c.oracle_voxel_f32(x[0].ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(3000), None, ctypes.c_float(0.5),
                                  cx.ctypes.data_as(ctypes.c_void_p), cc.ctypes.data_as(ctypes.c_void_p))
    cent, cnt, _ = osm.voxel_downsample(x[0], 0.5)
    assert v == cent.shape[0] and np.array_equal(cc[:v], cnt) and np.array_equal(cx[:v], cent)


def test_golden_fixtures_reproduce():
    """tests/golden/*.npz were written by tests/golden/make_golden.py from the oracle; they pin the
    oracle against accidental edits (and are what the GPU tests compare with on the box)."""
    from pathlib import Path
    g = np.load(Path(__file__).parent / "golden" / "pointnet_small.npz")
    p = onp.make_params(int(g["C"]), int(g["P"]), seed=int(g["weight_seed"]))
    x = synth.gaussian_clouds(int(g["input_seed"]), int(g["b"]), int(g["n"]))
    o = onp.forward(x, p)
    for k in ("cls", "seg", "R", "cls_logits", "seg_logits"):
        assert np.abs(o[k] - g[k]).max() <= 1e-5 * max(1.0, np.abs(g[k]).max()), k
    s = np.load(Path(__file__).parent / "golden" / "sampling_small.npz")
    pts = synth.lidar_scans(int(s["seed"]), 2, int(s["n"]))
    assert np.array_equal(osm.fps_batch(pts, int(s["s"])), s["fps_idx"])
    cent, cnt, keys = osm.voxel_downsample(pts[0], float(s["voxel"]))
    assert np.array_equal(cent, s["vox_cent"]) and np.array_equal(cnt, s["vox_cnt"])
