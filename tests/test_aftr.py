"""AftrBurner frame writer / parser — CPU only.  The writer is held byte for byte to files the REFERENCE's own
Provizio.to_aftr_frame wrote in this container (tests/golden/make_aftr_golden.py -> aftr_frames.npz): float32 and float64
points incl. signed zeros, the 1e-4 / 1e16 notation boundaries, denormals, inf / nan; with and without labels."""
import os
from pathlib import Path

import numpy as np
import pytest

from pointcloudprocessing_b200.pointcloud import aftr

G = np.load(Path(__file__).parent / "golden" / "aftr_frames.npz")


@pytest.mark.parametrize("tag", ["f32", "f64"])
@pytest.mark.parametrize("ltag", ["nolabels", "labels"])
def test_writer_reproduces_the_reference_files(lib, tag, ltag):
    pts = G[f"{tag}_points"]
    lab = G[f"{tag}_labels"] if ltag == "labels" else np.array([])
    want = G[f"{tag}_{ltag}_bytes"].tobytes()
    got = aftr.format_aftr_frame(pts, lab)
    if got != want:
        gl, wl = got.split(b"\n"), want.split(b"\n")
        first = next(i for i in range(min(len(gl), len(wl))) if gl[i] != wl[i])
        raise AssertionError((first, gl[first], wl[first]))


def test_to_aftr_frame_checks_and_round_trip(lib, tmp_path):
    msgs = []
    aftr.to_aftr_frame(str(tmp_path / "a.txt"), np.zeros((4, 2)), print_func=msgs.append)
    aftr.to_aftr_frame(str(tmp_path / "a.txt"), np.zeros((4, 3)), np.array(["x"] * 3), print_func=msgs.append)
    aftr.to_aftr_frame(str(tmp_path / "nodir" / "a.txt"), np.zeros((4, 3)), print_func=msgs.append)
    assert len(msgs) == 3 and "must be shape (N, 3)" in msgs[0] and "4 points and 3 labels" in msgs[1] and "path does not exist" in msgs[2]
    assert not os.path.exists(tmp_path / "a.txt")
    # writer -> parser round trip: float64 coordinates come back bit for bit (shortest round-trip digits), labels as indices
    rng = np.random.default_rng(3)
    classes, parts = ["kc46", "f15"], ["wing", "fuselage", "engine"]
    pts = rng.normal(0, 80, (5000, 3))
    pl = rng.integers(0, 3, 5000)
    lab = np.stack([np.array(["f15"] * 5000), np.array(parts)[pl]], 1)
    aftr.to_aftr_frame(str(tmp_path / "frame_0.txt"), pts, lab)
    xyz, cls, part, bad = aftr.parse_aftr_frame(open(tmp_path / "frame_0.txt", "rb").read(), classes, parts)
    assert bad == 0 and cls == 1 and np.array_equal(xyz, pts) and np.array_equal(part, pl)
    # float32 points: the text holds the digits of the float32's double value, which reads back to the same float32
    p32 = pts.astype(np.float32)
    xyz32, _, _, _ = aftr.parse_aftr_frame(aftr.format_aftr_frame(p32, lab), classes, parts)
    assert np.array_equal(xyz32.astype(np.float32), p32)
