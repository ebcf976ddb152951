"""Golden AftrBurner frame files written by the REFERENCE's own writer run here: Provizio.to_aftr_frame is lifted out of
/root/reference/point_cloud_toolkit/utils/Provizio.py with ast (the module itself imports the toolkit's GUI dependencies) and
executed on seeded float32 and float64 points, with and without labels.  Writes tests/golden/aftr_frames.npz (inputs + the
bytes the reference wrote).  Run from the repo root: python tests/golden/make_aftr_golden.py"""
import ast
import os
import sys
import tempfile
from pathlib import Path

import numpy as np

REF = Path("/root/reference/point_cloud_toolkit/utils/Provizio.py")
src = REF.read_text()
fn = next(n for n in ast.walk(ast.parse(src)) if isinstance(n, ast.FunctionDef) and n.name == "to_aftr_frame")
ns = {"np": np, "os": os}
exec(compile(ast.Module(body=[fn], type_ignores=[]), str(REF), "exec"), ns)


class Self:
    msgs = []
    def _print(self, m):
        self.msgs.append(m)


rng = np.random.default_rng(77)
special = np.array([[0.0, -0.0, 1.0], [1e-4, 9.999e-5, 1e16], [123456789.125, 1e15, 1.5e-7], [-2.5, 100000.0, 3.0],
                    [np.inf, -np.inf, np.nan], [1e-45, 3.4028235e38, 16777216.0]], np.float64)
cases = {}
for tag, dt in (("f32", np.float32), ("f64", np.float64)):
    pts = np.concatenate([special, np.round(rng.normal(0, 60, (200, 3)), 3), rng.normal(0, 1e-3, (20, 3)), rng.normal(0, 1e9, (20, 3))]).astype(dt)
    lab = np.stack([rng.choice(["kc46", "f15"], len(pts)), rng.choice(["wing", "fuselage", "engine", "boom"], len(pts))], 1)
    for ltag, l in (("nolabels", np.array([])), ("labels", lab)):
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "frame_0.txt")
            ns["to_aftr_frame"](Self(), path, pts, l)
            cases[f"{tag}_{ltag}_bytes"] = np.frombuffer(open(path, "rb").read(), np.uint8)
    cases[f"{tag}_points"] = pts
    cases[f"{tag}_labels"] = lab
out = Path(__file__).parent / "aftr_frames.npz"
np.savez_compressed(out, **cases)
print(out, {k: v.shape for k, v in cases.items()}, "numpy", np.__version__, file=sys.stderr)
