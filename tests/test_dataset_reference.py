"""CPU: the data-layer mirror against OUTPUTS OF THE REFERENCE ITSELF — tests/golden/dataset_reference.npz holds what the
unmodified reference class PointCloudSet (point_cloud_analysis/pointcloud/PointCloudSet.py) produced in this container
for a committed synthetic AftrBurner collect (generator: tests/golden/make_dataset_golden.py; TensorFlow replaced by a
protobuf stand-in there).  Parity pinned for: the frame text parser incl. which frames are rejected and why
(PointCloudSet.py:155-214), the pose log parser (472-509), _adjust_to_input_width (443-470), the aligned shuffle and the
test/val/train split arithmetic with class / part accounting (220-292), _serialize_sample (306-323), get_info (396-441).

The reference draws from NumPy's global legacy stream; the mirror owns a Generator.  For these tests the
mirror's stream is swapped for `np.random` itself, seeded as the generator seeded it, so every draw is the same draw."""
import json
from pathlib import Path

import numpy as np
import pytest

from pointcloudprocessing_b200.pointcloud import PointCloudSet, tfrecord

G = np.load(Path(__file__).parent / "golden" / "dataset_reference.npz")
CLASSES, PARTS, WIDTH, SEED = [str(c) for c in G["classes"]], [str(p) for p in G["parts"]], int(G["width"]), int(G["seed"])


class LegacyStream:
    """np.random's module-level stream behind the two calls the mirror makes (picklable, unlike the module)."""

    def uniform(self, lo, hi, size):
        return np.random.uniform(lo, hi, size)

    def shuffle(self, a):
        np.random.shuffle(a)


def write_collect(root: Path):
    d = root / "collect_2026.Nov.04_00.00.00.UTC"
    (d / "Lidar").mkdir(parents=True)
    for i, t in enumerate(G["frames"]):
        (d / "Lidar" / f"frame_{i}.txt").write_text(str(t))
    (d / "Lidar" / "_Lidar__2026.log").write_text("sensor log\n")
    (d / "_palindrome_state__2026.Nov.04.log").write_text(str(G["state_log"]))
    return d


def read_split(path: Path):
    f = tfrecord.TFRecordFile(str(path))
    rows = [f.decode(i, WIDTH) for i in range(len(f))]
    if not rows:
        return np.zeros((0, WIDTH * 3), np.float32), np.zeros((0,), np.int64), np.zeros((0, WIDTH), np.int64), np.zeros((0, 9), np.float32)
    return (np.stack([r[0].reshape(-1) for r in rows]), np.array([r[1] for r in rows], np.int64),
            np.stack([r[2].astype(np.int64) for r in rows]), np.stack([r[3].reshape(-1) for r in rows]))


@pytest.fixture()
def mirror(lib, tmp_path):
    msgs = []
    pcs = PointCloudSet("golden", CLASSES, PARTS, WIDTH, val=0.2, test=0.1, rand_seed=7, description="golden set",
                        print_func=lambda *a, **k: msgs.append(str(a[0]) if a else ""), data_path=str(tmp_path) + "/")
    pcs._rng = LegacyStream()                 # the reference's stream: np.random.uniform / np.random.shuffle
    return pcs, msgs, tmp_path


def test_pose_log_matches_reference(mirror):
    pcs, _, tmp = mirror
    d = write_collect(tmp)
    st = pcs._parse_state_info(str(d / "_palindrome_state__2026.Nov.04.log"))
    assert sorted(st.keys()) == list(G["state_frames"])
    for key in ("Sensor Pose", "Tanker Pose", "tanker_in_sensor_frame"):
        got = np.stack([st[k][key] for k in sorted(st.keys())])
        assert np.array_equal(got, G["state_" + key.replace(" ", "_")]), key          # same float64 arithmetic, bit for bit
    assert [st[k]["Time"] for k in sorted(st.keys())] == [str(t) for t in G["state_time"]]


def test_adjust_to_input_width_matches_reference(mirror):
    pcs, _, _ = mirror
    np.random.seed(SEED + 2)
    o, p = pcs._adjust_to_input_width(np.arange(15 * 3, dtype=np.float64).reshape(15, 3), np.arange(15) % 4)
    assert np.array_equal(o, G["adj_obs"]) and np.array_equal(p, G["adj_pl"])


def test_ingest_split_and_accounting_match_reference(mirror):
    pcs, msgs, tmp = mirror
    d = write_collect(tmp)
    np.random.seed(SEED)
    assert pcs.add_from_aftr_output(str(d), shuffle_points=True) is True
    for split in ("test", "val", "train"):
        o, c, s, e = read_split(tmp / "golden" / d.name / f"{split}_0.tfrecord")
        assert np.array_equal(o, G[f"a_{split}_obs"]), split                         # float32 bits of every stored coordinate
        assert np.array_equal(c, G[f"a_{split}_cls"]) and np.array_equal(s, G[f"a_{split}_seg"]) and np.array_equal(e, G[f"a_{split}_se3"]), split
    # the same frames are rejected, for the same stated reason (exception class names differ where the native parser raises)
    want = [str(m).replace("<DIR>", str(d)) for m in G["a_messages"]]
    assert len(msgs) == len(want)
    for got, ref in zip(msgs, want):
        assert got.split("\n")[0] == ref.split("\n")[0]
        if "\n" in ref:
            assert got.split(" : ", 1)[1] == ref.split(" : ", 1)[1], (got, ref)
    # second set through add_data, unshuffled: sizes accumulate across sets
    pcs.add_data("second", G["b_obs"], G["b_cls"], G["b_seg"], G["b_se3"], shuffle_points=False)
    for split in ("test", "val", "train"):
        o, c, s, e = read_split(tmp / "golden" / "second" / f"{split}_1.tfrecord")
        assert np.array_equal(o, G[f"b_{split}_obs"]) and np.array_equal(c, G[f"b_{split}_cls"]), split
        assert np.array_equal(s, G[f"b_{split}_seg"]) and np.array_equal(e, G[f"b_{split}_se3"]), split
    assert json.loads(json.dumps(pcs._data_size, default=int)) == json.loads(str(G["data_size_json"]))
    assert pcs.get_info() == str(G["info"])
