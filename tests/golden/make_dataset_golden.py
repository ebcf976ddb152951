"""Golden vectors for the data layer either side of the hot path (SURVEY 8 row a9 + 8f rank 1/2) from the REFERENCE ITSELF:
the unmodified class `PointCloudSet` of point_cloud_analysis/pointcloud/PointCloudSet.py is imported in this container
and driven through its own public calls (constructor, add_from_aftr_output, add_data, get_info).

TensorFlow is absent here, so a stand-in module named `tensorflow` is put in sys.modules before the import.  It supplies
only what that file touches: tf.train.{Example,Features,Feature,FloatList,Int64List} as google.protobuf classes built
from the published tensorflow/core/example/*.proto schema, a TFRecordWriter that records each payload it is handed, and
inert tf.io.FixedLenFeature / tf.float32 / tf.int64 placeholders.  Everything on the path under test — the frame text
parser, the pose-log parser, _adjust_to_input_width, the shuffle, the test/val/train split arithmetic, the class / part
accounting, _serialize_sample — is the reference's own code executing.

The reference draws from NumPy's global legacy RNG (np.random.uniform / np.random.shuffle); the generator seeds it with
np.random.seed(SEED) right before the ingest call and the parity test makes the mirror draw from the same stream.

Writes tests/golden/dataset_reference.npz (inputs as text + decoded outputs).  Run:
    python tests/golden/make_dataset_golden.py          (needs /root/reference)"""
import importlib.util
import json
import sys
import tempfile
import types
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
SEED = 20261104
WIDTH = 64
CLASSES = ["f-15_model", "a-10", "kc-46"]
PARTS = ["wing", "fuselage", "engine", "hstab"]
MESSAGES = []


def log(msg=""):
    MESSAGES.append(str(msg))


def example_classes():
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fd = descriptor_pb2.FileDescriptorProto(name="tf_example_golden.proto", package="tensorflow", syntax="proto3")
    T = descriptor_pb2.FieldDescriptorProto

    def msg(name):
        m = fd.message_type.add()
        m.name = name
        return m
    m = msg("BytesList"); m.field.add(name="value", number=1, type=T.TYPE_BYTES, label=T.LABEL_REPEATED)
    m = msg("FloatList"); f = m.field.add(name="value", number=1, type=T.TYPE_FLOAT, label=T.LABEL_REPEATED); f.options.packed = True
    m = msg("Int64List"); f = m.field.add(name="value", number=1, type=T.TYPE_INT64, label=T.LABEL_REPEATED); f.options.packed = True
    m = msg("Feature")
    m.oneof_decl.add(name="kind")
    m.field.add(name="bytes_list", number=1, type=T.TYPE_MESSAGE, type_name=".tensorflow.BytesList", label=T.LABEL_OPTIONAL, oneof_index=0)
    m.field.add(name="float_list", number=2, type=T.TYPE_MESSAGE, type_name=".tensorflow.FloatList", label=T.LABEL_OPTIONAL, oneof_index=0)
    m.field.add(name="int64_list", number=3, type=T.TYPE_MESSAGE, type_name=".tensorflow.Int64List", label=T.LABEL_OPTIONAL, oneof_index=0)
    m = msg("Features")
    e = m.nested_type.add(name="FeatureEntry")
    e.options.map_entry = True
    e.field.add(name="key", number=1, type=T.TYPE_STRING, label=T.LABEL_OPTIONAL)
    e.field.add(name="value", number=2, type=T.TYPE_MESSAGE, type_name=".tensorflow.Feature", label=T.LABEL_OPTIONAL)
    m.field.add(name="feature", number=1, type=T.TYPE_MESSAGE, type_name=".tensorflow.Features.FeatureEntry", label=T.LABEL_REPEATED)
    m = msg("Example")
    m.field.add(name="features", number=1, type=T.TYPE_MESSAGE, type_name=".tensorflow.Features", label=T.LABEL_OPTIONAL)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    get = lambda n: message_factory.GetMessageClass(pool.FindMessageTypeByName("tensorflow." + n))
    return {n: get(n) for n in ("Example", "Features", "Feature", "FloatList", "Int64List", "BytesList")}


WRITTEN = {}            # file name -> [payload bytes]


class RecordingWriter:
    def __init__(self, path):
        self.key = Path(path).name
        WRITTEN[self.key] = []

    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False

    def write(self, payload):
        WRITTEN[self.key].append(bytes(payload))


def install_tensorflow_stand_in():
    tf = types.ModuleType("tensorflow")
    tf.io = types.SimpleNamespace(FixedLenFeature=lambda shape, dtype: ("FixedLenFeature", tuple(shape), dtype), TFRecordWriter=RecordingWriter)
    tf.float32, tf.int64 = "float32", "int64"
    tf.train = types.SimpleNamespace(**example_classes())
    sys.modules["tensorflow"] = tf
    return tf


def load_reference_module():
    root = next(REF.rglob("point_cloud_analysis/pointcloud/PointCloudSet.py")).parent.parent
    sys.path.insert(0, str(root))
    spec = importlib.util.spec_from_file_location("ref_pointcloudset", root / "pointcloud" / "PointCloudSet.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_pointcloudset"] = mod
    spec.loader.exec_module(mod)
    return mod


def frame_text(rng, n, cls, bad=None):
    lines = []
    for j in range(n):
        x, y, z = rng.normal(0, 25, 3)
        part = PARTS[int(rng.integers(0, len(PARTS)))]
        sx = f"{x:.6f}"
        if bad == "nan" and j % 7 == 3:
            sx = "nan"
        if bad == "inf" and j % 9 == 1:
            sx = "-inf"
        sep = "  " if j % 5 == 0 else " "                  # the logger pads some columns; the reference drops the '' tokens
        lines.append(f"({sx}, {y:.6f}, {z:.4e}){sep}{cls}{sep}{part}")
    if bad == "part":
        lines[n // 2] = lines[n // 2].rsplit(" ", 1)[0] + " rotor"
    if bad == "class":
        lines[1] = lines[1].replace(cls, "zeppelin")
    if bad == "blank":
        lines.insert(n // 3, "")
    if bad == "number":
        lines[2] = lines[2].replace("(", "(1.2.3")
    if bad == "onelabel":
        lines[0] = lines[0].rsplit(" ", 1)[0]
    return "\n".join(lines) + ("\n" if n else "")


def make_inputs():
    rng = np.random.default_rng(SEED)
    plan = [(200, None), (64, None), (30, None), (1, None), (150, "nan"), (40, "inf"), (90, "part"), (0, None), (70, "blank"),
            (65, None), (63, None), (80, "class"), (100, "number"), (12, None), (300, None), (64, "onelabel"), (75, None), (20, "nan"),
            (128, None), (50, None), (64, None), (90, None), (10, None), (77, None)]
    frames = [frame_text(rng, n, CLASSES[i % 3], bad) for i, (n, bad) in enumerate(plan)]
    rows = ["Time   Frame   Sensor Pose   Tanker Pose"]
    for i in range(len(plan)):
        a, b = rng.uniform(0, 2 * np.pi, 2)
        S, T = np.eye(4), np.eye(4)
        S[:3, :3] = [[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]]
        T[:3, :3] = [[1, 0, 0], [0, np.cos(b), -np.sin(b)], [0, np.sin(b), np.cos(b)]]
        S[:3, 3], T[:3, 3] = rng.uniform(-5, 5, 3), rng.uniform(-50, 50, 3)
        rows.append(f"{i * 0.05:.3f} {i} " + " ".join(f"{v:.9f}" for v in S.T.ravel()) + " " + " ".join(f"{v:.9f}" for v in T.T.ravel()))
    return frames, "\n".join(rows) + "\n"


def write_collect(root: Path, frames, state_log):
    d = root / "collect_2026.Nov.04_00.00.00.UTC"
    (d / "Lidar").mkdir(parents=True)
    for i, t in enumerate(frames):
        (d / "Lidar" / f"frame_{i}.txt").write_text(t)
    (d / "Lidar" / "_Lidar__2026.log").write_text("sensor log\n")     # listdir is one longer than the frames: frame_N is probed and fails
    (d / "_palindrome_state__2026.Nov.04.log").write_text(state_log)
    return d


def decode(payloads, Example):
    obs, cls, seg, se3 = [], [], [], []
    for p in payloads:
        ex = Example.FromString(p)
        f = ex.features.feature
        obs.append(np.array(f["observations"].float_list.value, np.float32))
        cls.append(int(f["class_label"].int64_list.value[0]))
        seg.append(np.array(f["part_labels"].int64_list.value, np.int64))
        se3.append(np.array(f["se3"].float_list.value, np.float32))
    z = lambda a, w, dt: np.stack(a) if a else np.zeros((0, w), dt)
    return z(obs, WIDTH * 3, np.float32), np.array(cls, np.int64), z(seg, WIDTH, np.int64), z(se3, 9, np.float32)


def main():
    tf = install_tensorflow_stand_in()
    ref = load_reference_module()
    frames, state_log = make_inputs()
    out = {"frames": np.array(frames), "state_log": np.array(state_log), "seed": SEED, "width": WIDTH,
           "classes": np.array(CLASSES), "parts": np.array(PARTS)}
    with tempfile.TemporaryDirectory() as tmp:
        tmp = Path(tmp)
        d = write_collect(tmp, frames, state_log)
        pcs = ref.PointCloudSet("golden", CLASSES, PARTS, WIDTH, val=0.2, test=0.1, rand_seed=7, description="golden set",
                                print_func=log, data_path=str(tmp) + "/")
        # pose log on its own (PointCloudSet.py:472-509)
        state = pcs._parse_state_info(str(d / "_palindrome_state__2026.Nov.04.log"))
        out["state_frames"] = np.array(sorted(state.keys()))
        for key in ("Sensor Pose", "Tanker Pose", "tanker_in_sensor_frame"):
            out["state_" + key.replace(" ", "_")] = np.stack([state[k][key] for k in sorted(state.keys())])
        out["state_time"] = np.array([state[k]["Time"] for k in sorted(state.keys())])
        # ingest, shuffled, from the seeded global stream
        np.random.seed(SEED)
        assert pcs.add_from_aftr_output(str(d), shuffle_points=True) is True
        for split in ("test", "val", "train"):
            o, c, s, e = decode(WRITTEN[f"{split}_0.tfrecord"], tf.train.Example)
            out[f"a_{split}_obs"], out[f"a_{split}_cls"], out[f"a_{split}_seg"], out[f"a_{split}_se3"] = o, c, s, e
        out["a_messages"] = np.array([m.replace(str(d), "<DIR>") for m in MESSAGES])
        # a second set through add_data, unshuffled (PointCloudSet.py:220-292): split arithmetic + accounting accumulate
        rng = np.random.default_rng(SEED + 1)
        N = 23
        obs = rng.normal(0, 10, (N, WIDTH, 3))
        cl = rng.integers(0, len(CLASSES), N)
        pl = rng.integers(0, len(PARTS), (N, WIDTH))
        se = rng.normal(0, 1, (N, 3, 3))
        pcs.add_data("second", obs, cl, pl, se, shuffle_points=False)
        out["b_obs"], out["b_cls"], out["b_seg"], out["b_se3"] = obs, cl, pl, se
        for split in ("test", "val", "train"):
            o, c, s, e = decode(WRITTEN[f"{split}_1.tfrecord"], tf.train.Example)
            out[f"b_{split}_obs"], out[f"b_{split}_cls"], out[f"b_{split}_seg"], out[f"b_{split}_se3"] = o, c, s, e
        sizes = json.loads(json.dumps(pcs._data_size, default=int))
        out["data_size_json"] = np.array(json.dumps(sizes, sort_keys=True))
        out["info"] = np.array(pcs.get_info())
        # _adjust_to_input_width alone (PointCloudSet.py:443-470), global stream re-seeded
        np.random.seed(SEED + 2)
        small = np.arange(15 * 3, dtype=np.float64).reshape(15, 3)
        o, p = pcs._adjust_to_input_width(small, np.arange(15) % 4)
        out["adj_obs"], out["adj_pl"] = o, p
    dst = Path(__file__).with_name("dataset_reference.npz")
    np.savez_compressed(dst, **out)
    print("wrote", dst, {k: getattr(v, "shape", None) for k, v in out.items() if k.startswith(("a_", "b_t", "b_v"))})
    print("\n".join(out["a_messages"]))
    print(out["info"])


if __name__ == "__main__":
    main()
