"""Golden vectors for the pose post-processing row from the REFERENCE ITSELF, run in this container:
  PCA/utils/mat_ops.py  : get_dcm, get_roll_pitch_yaw_deg, reorthogonalize          (imported; plotly is stubbed — unused)
  PCT/utils/calibration.py : solve_kabsch   (its module does `from dependencies import *` (PyQt ...), so the function's own
                             source is compiled from the reference file with numpy in scope — nothing is copied here)
Writes tests/golden/pose_reference.npz.  Run:  python tests/golden/make_pose_golden.py   (needs /root/reference)."""
import ast
import importlib.util
import sys
import types
from pathlib import Path

import numpy as np

REF = Path("/root/reference")


def load_mat_ops():
    for name in ("plotly", "plotly.graph_objects", "plotly.subplots"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["plotly.subplots"].make_subplots = lambda *a, **k: None
    path = next(REF.rglob("point_cloud_analysis/utils/mat_ops.py"))
    spec = importlib.util.spec_from_file_location("ref_mat_ops", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def load_solve_kabsch():
    path = next(REF.rglob("utils/calibration.py"))
    tree = ast.parse(path.read_text())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "solve_kabsch")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), str(path), "exec"), ns)
    return ns["solve_kabsch"]


def main():
    mo, kabsch = load_mat_ops(), load_solve_kabsch()
    rng = np.random.default_rng(20261104)
    rpy = np.concatenate([rng.uniform(-180, 180, (60, 3)), [[0, 0, 0], [0, 0, 90], [10, -89.5, 30], [180, 45, -180], [-33, 89.9, 12]]])
    dcm = np.stack([mo.get_dcm(*r) for r in rpy])
    back = np.array([[d[k] for k in ("roll", "pitch", "yaw")] for d in (mo.get_roll_pitch_yaw_deg(m) for m in dcm)])
    noisy = dcm + rng.normal(0, 0.05, dcm.shape)
    noisy[3] = -noisy[3]                                   # a reflection: reorthogonalize keeps det < 0 (no sign fix in the reference)
    noisy = np.concatenate([noisy, rng.normal(0, 1, (20, 3, 3))])
    reo = np.stack([mo.reorthogonalize(m) for m in noisy])
    # Kabsch: measured = R^T (truth - t) + noise, three sizes, one near-planar set and one mirrored set (det fix branch)
    ks = []
    for i, n in enumerate([4, 12, 100, 1000, 50, 30]):
        truth = rng.normal(0, 5, (n, 3))
        if i == 4:
            truth[:, 2] *= 1e-3
        R = mo.get_dcm(*rng.uniform(-180, 180, 3))
        t = rng.normal(0, 2, 3)
        meas = (truth - t) @ R + rng.normal(0, 0.02, (n, 3))
        if i == 5:
            meas[:, 0] *= -1.0
        Rk, tk, rmse = kabsch(truth, meas)
        ks.append((truth.astype(np.float32), meas.astype(np.float32)))
    kout = [kabsch(t.astype(np.float64), m.astype(np.float64)) for t, m in ks]          # on the float32-rounded inputs the device sees
    np.savez(Path(__file__).parent / "pose_reference.npz", rpy=rpy, dcm=dcm, rpy_back=back, noisy=noisy.astype(np.float32),
             reorth=np.stack([mo.reorthogonalize(m.astype(np.float64)) for m in noisy.astype(np.float32)]), reorth_f64_in=reo,
             **{f"k{i}_truth": ks[i][0] for i in range(len(ks))}, **{f"k{i}_meas": ks[i][1] for i in range(len(ks))},
             **{f"k{i}_R": kout[i][0] for i in range(len(ks))}, **{f"k{i}_t": kout[i][1] for i in range(len(ks))},
             **{f"k{i}_rmse": np.float64(kout[i][2]) for i in range(len(ks))}, n_kabsch=np.int64(len(ks)))
    print("wrote pose_reference.npz:", len(rpy), "dcms,", len(noisy), "reorthogonalisations,", len(ks), "Kabsch problems")


if __name__ == "__main__":
    main()
