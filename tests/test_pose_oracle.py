"""CPU: the pose oracle (oracle/pose.py) against outputs of the REFERENCE's own functions (PCA/utils/mat_ops.py,
PCT/utils/calibration.py) recorded in tests/golden/pose_reference.npz — this oracle row is parity-pinned."""
from pathlib import Path

import numpy as np

from oracle import pose as op

G = np.load(Path(__file__).parent / "golden" / "pose_reference.npz")


def test_dcm_and_angles_match_reference():
    for rpy, dcm, back in zip(G["rpy"], G["dcm"], G["rpy_back"]):
        assert np.allclose(op.get_dcm(*rpy), dcm, rtol=0, atol=1e-15)
        assert np.allclose(op.get_roll_pitch_yaw_deg(dcm), back, rtol=0, atol=1e-12)


def test_reorthogonalize_matches_reference():
    for m, r in zip(G["noisy"], G["reorth"]):
        assert np.allclose(op.reorthogonalize(m), r, rtol=0, atol=1e-13)
    assert np.linalg.det(G["reorth"][3]) < 0                                  # the reference keeps reflections


def test_kabsch_matches_reference():
    for i in range(int(G["n_kabsch"])):
        R, t, rmse = op.solve_kabsch(G[f"k{i}_truth"], G[f"k{i}_meas"])
        assert np.allclose(R, G[f"k{i}_R"], atol=1e-12) and np.allclose(t, G[f"k{i}_t"], atol=1e-11)
        assert abs(rmse - float(G[f"k{i}_rmse"])) < 1e-11
        assert np.linalg.det(R) > 0
