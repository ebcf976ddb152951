"""CPU: the data layer either side of the hot path — TFRecord / tf.train.Example I/O and the AftrBurner frame parser
(native, csrc/host_io.cu) and the PointCloudSet mirror (pointcloud/PointCloudSet.py of the reference).

Pins: CRC32C known-answer vectors (RFC 3720 B.4), the Example wire format cross-checked against google.protobuf's own
serializer on dynamically built tensorflow.Example descriptors (an independent implementation), and a line-by-line
Python restatement of the reference's frame parsing loop (PointCloudSet.py:161-198)."""
import ctypes as C
import os
import struct

import numpy as np
import pytest

from pointcloudprocessing_b200 import _ffi
from pointcloudprocessing_b200.pointcloud import PointCloudSet, load_from_file, tfrecord

CLASSES = ["f-15_model", "a-10", "kc-46"]
PARTS = ["wing", "fuselage", "engine", "hstab"]


def test_crc32c_known_answers(lib):
    assert lib.pcp_crc32c(b"123456789", 9) == 0xE3069283
    assert lib.pcp_crc32c(bytes(32), 32) == 0x8A9136AA                    # RFC 3720 B.4: 32 bytes of zeros
    assert lib.pcp_crc32c(bytes([0xFF] * 32), 32) == 0x62A8AB43            # 32 bytes of ones
    assert lib.pcp_crc32c(bytes(range(32)), 32) == 0x46DD794E              # incrementing
    crc = lib.pcp_crc32c(b"abc", 3)
    assert lib.pcp_crc32c_masked(b"abc", 3) == ((((crc >> 15) | (crc << 17)) + 0xA282EAD8) & 0xFFFFFFFF)


def _example_classes():
    """tensorflow/core/example/{feature,example}.proto rebuilt with google.protobuf (no TensorFlow needed)."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fd = descriptor_pb2.FileDescriptorProto(name="tf_example_test.proto", package="tensorflow", syntax="proto3")
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
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("tensorflow.Example"))


def _sample(rng, width, big_labels=False):
    obs = rng.standard_normal((width, 3)).astype(np.float32) * 30
    seg = rng.integers(0, 300 if big_labels else len(PARTS), width).astype(np.int64)
    se3 = rng.standard_normal((3, 3)).astype(np.float32)
    return obs, int(rng.integers(0, len(CLASSES))), seg, se3


@pytest.mark.parametrize("big_labels", [False, True])
def test_example_wire_format_matches_protobuf(lib, big_labels):
    Example = _example_classes()
    rng = np.random.default_rng(0)
    width = 37
    obs, cl, seg, se3 = _sample(rng, width, big_labels)
    mine = tfrecord.serialize_sample(obs, cl, seg, se3)
    ex = Example()
    ex.features.feature["observations"].float_list.value.extend(obs.flatten().tolist())
    ex.features.feature["class_label"].int64_list.value.append(cl)
    ex.features.feature["part_labels"].int64_list.value.extend(seg.tolist())
    ex.features.feature["se3"].float_list.value.extend(se3.flatten().tolist())
    theirs = ex.SerializeToString(deterministic=True)
    assert mine == theirs                                        # byte-identical to protobuf's deterministic serializer
    # protobuf parses ours
    back = Example.FromString(mine)
    assert list(back.features.feature["part_labels"].int64_list.value) == seg.tolist()
    # the native decoder parses theirs (any map order) and ours
    for blob in (theirs, ex.SerializeToString(), mine):
        o = np.empty((width, 3), np.float32); p = np.empty(width, np.int32); s = np.empty((3, 3), np.float32); c = C.c_int64()
        _ffi.check(lib.pcp_example_decode(blob, len(blob), width, _ffi.ptr(o), C.byref(c), _ffi.ptr(p), _ffi.ptr(s)))
        assert np.array_equal(o, obs) and c.value == cl and np.array_equal(p, seg) and np.array_equal(s, se3)


def test_example_decode_accepts_unpacked_and_rejects_bad(lib):
    width = 3
    # hand-built unpacked encoding: FloatList value as repeated fixed32 (wire type 5), Int64List as repeated varint (wire type 0)
    def ld(field, payload):
        return tfrecord._varint((field << 3) | 2) + tfrecord._varint(len(payload)) + payload
    fl = lambda vals: ld(2, b"".join(b"\x0d" + struct.pack("<f", v) for v in vals))
    il = lambda vals: ld(3, b"".join(b"\x08" + tfrecord._varint(v) for v in vals))
    feats = {"observations": fl([float(i) for i in range(9)]), "class_label": il([2]), "part_labels": il([1, 200, 3]), "se3": fl([0.5] * 9)}
    blob = ld(1, b"".join(ld(1, ld(1, k.encode()) + ld(2, v)) for k, v in feats.items()))
    o = np.empty((width, 3), np.float32); p = np.empty(width, np.int32); s = np.empty((3, 3), np.float32); c = C.c_int64()
    _ffi.check(lib.pcp_example_decode(blob, len(blob), width, _ffi.ptr(o), C.byref(c), _ffi.ptr(p), _ffi.ptr(s)))
    assert o.ravel().tolist() == list(range(9)) and c.value == 2 and p.tolist() == [1, 200, 3] and float(s[0, 0]) == 0.5
    assert lib.pcp_example_decode(blob, len(blob), 4, _ffi.ptr(np.empty((4, 3), np.float32)), C.byref(c), _ffi.ptr(np.empty(4, np.int32)), _ffi.ptr(s)) == -1
    assert b"expected" in lib.pcp_last_error()
    assert lib.pcp_example_decode(blob[:-5], len(blob) - 5, width, _ffi.ptr(o), C.byref(c), _ffi.ptr(p), _ffi.ptr(s)) == -1
    del feats["se3"]
    blob2 = ld(1, b"".join(ld(1, ld(1, k.encode()) + ld(2, v)) for k, v in feats.items()))
    assert lib.pcp_example_decode(blob2, len(blob2), width, _ffi.ptr(o), C.byref(c), _ffi.ptr(p), _ffi.ptr(s)) == -1
    assert b"missing feature" in lib.pcp_last_error()


def test_tfrecord_roundtrip_and_corruption(lib, tmp_path):
    rng = np.random.default_rng(1)
    width = 64
    samples = [_sample(rng, width) for _ in range(7)]
    path = tmp_path / "t.tfrecord"
    with tfrecord.TFRecordWriter(str(path)) as w:
        for s in samples:
            w.write(tfrecord.serialize_sample(*s))
    f = tfrecord.TFRecordFile(str(path))
    assert len(f) == 7
    for i, (obs, cl, seg, se3) in enumerate(samples):
        o, c, p, s = f.decode(i, width)
        assert np.array_equal(o, obs) and c == cl and np.array_equal(p, seg) and np.array_equal(s, se3)
    # framing: uint64 length + masked crc of the length
    raw = path.read_bytes()
    n0 = struct.unpack("<Q", raw[:8])[0]
    assert n0 == len(tfrecord.serialize_sample(*samples[0]))
    assert struct.unpack("<I", raw[8:12])[0] == lib.pcp_crc32c_masked(raw[:8], 8)
    bad = bytearray(raw); bad[40] ^= 0x01
    (tmp_path / "bad.tfrecord").write_bytes(bytes(bad))
    with pytest.raises(_ffi.PcpError, match="corrupt payload CRC"):
        tfrecord.TFRecordFile(str(tmp_path / "bad.tfrecord"))
    assert len(tfrecord.TFRecordFile(str(tmp_path / "bad.tfrecord"), verify_crc=0)) == 7
    (tmp_path / "trunc.tfrecord").write_bytes(raw[:-3])
    with pytest.raises(_ffi.PcpError, match="truncated"):
        tfrecord.TFRecordFile(str(tmp_path / "trunc.tfrecord"))
    (tmp_path / "empty.tfrecord").write_bytes(b"")
    assert len(tfrecord.TFRecordFile(str(tmp_path / "empty.tfrecord"))) == 0


def reference_parse_lines(text, class_labels, part_labels):
    """Restatement of the per-line loop of PointCloudSet.add_from_aftr_output (PointCloudSet.py:161-198)."""
    obs, pl, cl, non_num = [], [], None, 0
    for line in text.splitlines():
        line = line.strip()
        a, b = line.find('('), line.find(')')
        pos = [float(v) for v in line[a + 1:b].split(',')]
        labels = [l for l in line[b + 1:].split(" ") if len(l) > 1]
        if len(labels) == 2:
            if labels[0] not in class_labels:
                raise Exception(f"Class label {labels[0]} not known")
            if labels[1] not in part_labels:
                raise Exception(f"Part label {labels[1]} not known")
        else:
            raise Exception("Dataset must contain both a class label and part label.")
        if np.isfinite(np.array(pos)).all():
            obs.append(np.array(pos)); cl = class_labels[labels[0]]; pl.append(part_labels[labels[1]])
        else:
            non_num += 1
    return np.array(obs), cl, np.array(pl), non_num


def make_frame_text(rng, n, cls="kc-46", nan_every=0):
    lines = []
    for i in range(n):
        x, y, z = rng.uniform(-40, 40, 3)
        if nan_every and i % nan_every == 3:
            lines.append(f"({x:.3f}, nan, {z:.3f}) {cls} {PARTS[i % len(PARTS)]}")
        elif nan_every and i % nan_every == 5:
            lines.append(f"(inf, {y:.3f}, {z:.3f}) {cls} {PARTS[i % len(PARTS)]}")
        else:
            lines.append(f"({x:.3f}, {y:.3f}, {z:.3f}) {cls} {PARTS[i % len(PARTS)]}")
    return "\n".join(lines) + "\n"


def make_set(tmp_path, width=128, **kw):
    return PointCloudSet("unit", CLASSES, PARTS, width, data_path=str(tmp_path) + "/", print_func=lambda *a, **k: None, **kw)


def test_frame_parser_matches_reference_restatement(lib, tmp_path):
    rng = np.random.default_rng(2)
    pcs = make_set(tmp_path)
    cl_map, pl_map = pcs._class_labels, pcs._part_labels
    for n, nan_every in ((490, 0), (313, 7), (1, 0)):
        text = make_frame_text(rng, n, nan_every=nan_every)
        p = tmp_path / f"frame_{n}.txt"
        p.write_text(text)
        obs, cl, pl, bad = pcs._parse_frame_file(str(p))
        r_obs, r_cl, r_pl, r_bad = reference_parse_lines(text, cl_map, pl_map)
        assert np.array_equal(obs, r_obs) and cl == r_cl and np.array_equal(pl, r_pl) and bad == r_bad
    # windows line ends, extra blanks between labels, scientific notation — all handled by strip()/split()/float()
    text = "(1e1, -2.5E-1,  3) kc-46  engine\r\n( 4.0,5.0,6.0 )   a-10 wing\r\n"
    (tmp_path / "w.txt").write_text(text)
    obs, cl, pl, bad = pcs._parse_frame_file(str(tmp_path / "w.txt"))
    r_obs, r_cl, r_pl, _ = reference_parse_lines(text, cl_map, pl_map)
    assert np.array_equal(obs, r_obs) and cl == r_cl and np.array_equal(pl, r_pl)
    # rejections: the reference raises inside its per-frame try block and skips the frame
    for bad_text, msg in (("(1,2,3) kc-46 rotor\n", "Part label rotor not known"), ("(1,2,3) zeppelin wing\n", "Class label zeppelin not known"),
                          ("(1,2,3) kc-46\n", "both a class label and part label"), ("(1,2,x) kc-46 wing\n", "could not convert"),
                          ("(1,2,3) kc-46 wing\n\n(1,2,3) kc-46 wing\n", "could not convert")):
        (tmp_path / "b.txt").write_text(bad_text)
        with pytest.raises(Exception, match=msg):
            pcs._parse_frame_file(str(tmp_path / "b.txt"))
        with pytest.raises(Exception):
            reference_parse_lines(bad_text, cl_map, pl_map)


def write_collect(tmp_path, rng, nframes, pts=(90, 200)):
    d = tmp_path / "collect_2026.Jan.22_00.00.00.UTC"
    (d / "Lidar").mkdir(parents=True)
    poses = []
    lines = ["Time   Frame   Sensor Pose   Tanker Pose"]
    for i in range(nframes):
        n = int(rng.integers(*pts))
        (d / "Lidar" / f"frame_{i}.txt").write_text(make_frame_text(rng, n, cls=CLASSES[i % 3], nan_every=11 if i == 1 else 0))
        S, T = np.eye(4), np.eye(4)
        a = rng.uniform(0, 2 * np.pi)
        S[:3, :3] = [[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]]
        S[:3, 3] = rng.uniform(-5, 5, 3); T[:3, 3] = rng.uniform(-50, 50, 3)
        poses.append((S, T))
        # 16 numbers per matrix, column-major (PointCloudSet.py:495-499 reads columns then transposes)
        lines.append(f"{i * 0.1:.3f} {i} " + " ".join(f"{v:.6f}" for v in S.T.ravel()) + " " + " ".join(f"{v:.6f}" for v in T.T.ravel()))
    (d / "Lidar" / "frame_x.txt").write_text("junk")                       # listdir length > number of real frames, like the reference logs
    (d / "_palindrome_state__2026.log").write_text("\n".join(lines) + "\n")
    return d, poses


def test_pointcloudset_end_to_end_on_cpu(lib, tmp_path):
    rng = np.random.default_rng(3)
    d, poses = write_collect(tmp_path, rng, 23)
    msgs = []
    pcs = PointCloudSet("kc46_unit", CLASSES, PARTS, 128, val=0.15, test=0.10, batch_size=4, rand_seed=5,
                        print_func=lambda *a, **k: msgs.append(a[0]), data_path=str(tmp_path) + "/")
    assert pcs.add_from_aftr_output(str(d))
    assert any("frame_23.txt" in m and "FileNotFoundError" in m for m in msgs)        # the reference's frame_5000-style log line
    sizes = {k: v["count"] for k, v in pcs._data_size.items()}
    assert sizes == {"test": 3, "val": 4, "train": 16}                                  # ceil(23*.1), ceil(23*.15), rest
    for split in ("train", "val", "test"):
        assert os.path.isfile(tmp_path / "kc46_unit" / d.name / f"{split}_0.tfrecord")
        assert sum(pcs._data_size[split]["part_count"].values()) == sizes[split] * 128    # every stored cloud is exactly input_width long
    info = pcs.get_info()
    assert "Total count: 16" in info and "Random seed: 5" in info
    # pose: tanker_in_sensor_frame = S_R^T T_R  | S_R^T (t_T - t_S)
    st = pcs._parse_state_info(str(d / "_palindrome_state__2026.log"))
    S, T = poses[7]
    assert np.allclose(st[7]["tanker_in_sensor_frame"][:3, :3], S[:3, :3].T @ T[:3, :3], atol=1e-5)
    assert np.allclose(st[7]["tanker_in_sensor_frame"][:3, 3], S[:3, :3].T @ (T[:3, 3] - S[:3, 3]), atol=1e-4)
    # iterator: infinite, batches of 4, every train sample seen once per pass, labels aligned with points
    it = pcs.get_train_set()
    seen = []
    for _ in range(8):                                                                    # two passes over 16 samples
        x, y = next(it)
        assert x.shape == (4, 128, 3) and x.dtype == np.float32
        assert y["classification_output"].shape == (4,) and y["segmentation_output"].shape == (4, 128) and y["se3"].shape == (4, 3, 3)
        assert np.isfinite(x).all() and y["segmentation_output"].max() < len(PARTS)
        seen += [float(v) for v in x[:, 0, 0]]
    assert len(set(seen[:16])) == 16 and sorted(seen[:16]) == sorted(seen[16:])
    # repeat() comes before batch(): batches are always full, spanning passes (3 test samples -> batch of 4)
    assert next(pcs.get_val_set())[0].shape == (4, 128, 3) and next(pcs.get_test_set())[0].shape == (4, 128, 3)
    assert [b[0].shape[0] for b in pcs.get_test_set(repeat=False)] == [3]
    # persistence
    again = load_from_file(str(tmp_path / "kc46_unit" / "pc_set.joblib"))
    assert again._data_size == pcs._data_size and again._network_input_width == 128
    assert load_from_file(str(tmp_path / "kc46_unit" / "pc_set"))._name == "kc46_unit"
    # jitter needs the GPU
    pj = PointCloudSet("kc46_unit", CLASSES, PARTS, 128, jitter_stdev_m=np.array([0.1, 0.1, 0.1]), batch_size=4, data_path=str(tmp_path) + "/",
                       print_func=lambda *a, **k: None)
    with pytest.raises(RuntimeError, match="GPU"):
        next(pj.get_train_set())


def test_adjust_to_input_width_semantics(lib, tmp_path):
    pcs = make_set(tmp_path, width=8, rand_seed=1)
    obs = np.arange(36, dtype=np.float64).reshape(12, 3)
    pl = np.arange(12)
    o, p = pcs._adjust_to_input_width(obs, pl)
    assert np.array_equal(o, obs[:8]) and np.array_equal(p, pl[:8])                       # truncate to the FIRST n (PointCloudSet.py:455-456)
    o, p = pcs._adjust_to_input_width(obs[:5], pl[:5])
    assert o.shape == (8, 3) and np.array_equal(o[:5], obs[:5])
    for k in range(5, 8):                                                                 # duplicates stay aligned with their labels
        assert p[k] in pl[:5] and np.array_equal(o[k], obs[p[k]])
    o, p = pcs._adjust_to_input_width(obs[:8], pl[:8])
    assert np.array_equal(o, obs[:8])
