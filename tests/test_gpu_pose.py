"""GPU: batched pose post-processing (csrc/pose.cu) against outputs of the REFERENCE's own functions
(tests/golden/pose_reference.npz, generated from PCA/utils/mat_ops.py and PCT/utils/calibration.py) and against the oracle
on random batches.  fp32 in / out around fp64 arithmetic: tolerances are fp32 round-off of the inputs / outputs."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import pose as op

pytestmark = pytest.mark.gpu
G = np.load(Path(__file__).parent / "golden" / "pose_reference.npz")


def test_dcm_and_angles_match_reference(lib):
    from pointcloudprocessing_b200 import ops
    dcm = ops.get_dcm(G["rpy"]).cpu().numpy()
    assert np.abs(dcm - G["dcm"]).max() < 2e-6                       # degrees arrive as fp32: 180 * 2^-24 deg = 1.9e-7 rad
    back = ops.get_roll_pitch_yaw_deg(G["dcm"]).cpu().numpy()
    # d(pitch)/d(dcm[0][2]) = 1 / cos(pitch): the fp32-rounded matrix entry dominates near the gimbal poles
    tol = 1e-4 + 6e-6 / np.maximum(np.cos(np.deg2rad(G["rpy_back"][:, 1])), 1e-3)
    assert (np.abs(back - G["rpy_back"]).max(axis=1) < tol).all()


def test_reorthogonalize_matches_reference(lib):
    from pointcloudprocessing_b200 import ops
    out = ops.reorthogonalize(G["noisy"]).cpu().numpy()
    assert np.abs(out - G["reorth"]).max() < 2e-6
    assert np.linalg.det(out[3].astype(np.float64)) < 0               # reflections are kept, like the reference
    rng = np.random.default_rng(3)
    m = rng.normal(0, 1, (4096, 3, 3)).astype(np.float32)
    m[::7] *= 1e-3
    out = ops.reorthogonalize(m).cpu().numpy().astype(np.float64)
    ref = np.stack([op.reorthogonalize(x) for x in m])
    cond = np.array([np.linalg.cond(x.astype(np.float64)) for x in m])
    ok = cond < 1e3                                                   # U V^T is ill-conditioned for near-singular inputs
    assert np.abs(out[ok] - ref[ok]).max() < 5e-5
    assert np.abs(np.einsum("bij,bkj->bik", out, out) - np.eye(3)).max() < 1e-6


def test_kabsch_matches_reference(lib):
    from pointcloudprocessing_b200 import ops
    for i in range(int(G["n_kabsch"])):
        R, t, rmse = ops.solve_kabsch(G[f"k{i}_truth"][None], G[f"k{i}_meas"][None])
        assert np.abs(R[0].cpu().numpy() - G[f"k{i}_R"]).max() < 2e-6, i
        assert np.abs(t[0].cpu().numpy() - G[f"k{i}_t"]).max() < 2e-5, i
        assert abs(float(rmse[0]) - float(G[f"k{i}_rmse"])) < 1e-5 * max(1.0, float(G[f"k{i}_rmse"])), i
    rng = np.random.default_rng(8)
    b, n = 64, 500
    truth = rng.normal(0, 3, (b, n, 3)).astype(np.float32)
    Rt = np.stack([op.get_dcm(*rng.uniform(-180, 180, 3)) for _ in range(b)])
    meas = (np.einsum("bnj,bjk->bnk", truth - 1.5, Rt) + rng.normal(0, 0.01, (b, n, 3))).astype(np.float32)
    R, t, rmse = ops.solve_kabsch(truth, meas)
    for c in range(b):
        Rr, tr, er = op.solve_kabsch(truth[c], meas[c])
        assert np.abs(R[c].cpu().numpy() - Rr).max() < 2e-6 and np.abs(t[c].cpu().numpy() - tr).max() < 2e-5
        assert abs(float(rmse[c]) - er) < 1e-6
