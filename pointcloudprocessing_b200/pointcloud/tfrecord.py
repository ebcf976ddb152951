"""TFRecord + tf.train.Example I/O without TensorFlow — the on-disk format of the reference's PointCloudSet
(pointcloud/PointCloudSet.py:251-288 `tf.io.TFRecordWriter`, :306-323 `_serialize_sample`, schema :100-105).

Record framing: uint64 length | uint32 masked_crc32c(length) | payload | uint32 masked_crc32c(payload).
Payload: Example{features=1: Features{feature=1: map<string, Feature>}}, Feature{float_list=2 | int64_list=3}, both
lists `repeated value = 1 [packed]`.  CRC, record scan and Example decode run natively in libpcp_b200.so
(csrc/host_io.cu); the encoder builds the bytes with NumPy (float lists are a memcpy, small-int lists one byte each).
"""
from __future__ import annotations

import ctypes as C
import struct

import os

import numpy as np

from .. import _ffi


def _varint(v: int) -> bytes:
    v &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _ld(field: int, payload: bytes) -> bytes:
    """length-delimited field"""
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def _varints_packed(values: np.ndarray) -> bytes:
    v = np.asarray(values, np.int64).ravel()
    if v.size and v.min() >= 0 and v.max() < 128:
        return v.astype(np.uint8).tobytes()                     # one byte per value (every label index in practice)
    return b"".join(_varint(int(x)) for x in v)


def _float_feature(values) -> bytes:
    raw = np.ascontiguousarray(np.asarray(values, "<f4").ravel()).tobytes()
    return _ld(2, _ld(1, raw))                                   # Feature.float_list -> FloatList.value (packed)


def _int64_feature(values) -> bytes:
    return _ld(3, _ld(1, _varints_packed(np.atleast_1d(values))))  # Feature.int64_list -> Int64List.value (packed)


def serialize_sample(obs, cls, seg, se3) -> bytes:
    """PointCloudSet._serialize_sample (PointCloudSet.py:306-323).  Map entries are emitted in sorted key order, as the
    protobuf C++ deterministic serializer TensorFlow uses does."""
    feats = {"class_label": _int64_feature(int(cls)), "observations": _float_feature(obs),
             "part_labels": _int64_feature(seg), "se3": _float_feature(se3)}
    body = b"".join(_ld(1, _ld(1, k.encode()) + _ld(2, v)) for k, v in sorted(feats.items()))
    return _ld(1, body)


class TFRecordWriter:
    """tf.io.TFRecordWriter subset: context manager with write(bytes)."""

    def __init__(self, path: str):
        self._f = open(path, "wb")
        self._lib = _ffi.load()

    def write(self, record: bytes):
        hdr = struct.pack("<Q", len(record))
        self._f.write(hdr)
        self._f.write(struct.pack("<I", self._lib.pcp_crc32c_masked(hdr, 8)))
        self._f.write(record)
        self._f.write(struct.pack("<I", self._lib.pcp_crc32c_masked(record, len(record))))

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class TFRecordFile:
    """A TFRecord file mapped into memory and indexed natively; `decode(i, width)` returns one PointCloudSet sample."""

    def __init__(self, path: str, verify_crc: int = 2):
        self._lib = _ffi.load()
        self.path = path
        # mapped, not read: a kc46-sized split is ~1.5 GB per process, and every rank opens every file
        self.buf = np.memmap(path, np.uint8, 'r') if os.path.getsize(path) > 0 else np.empty(0, np.uint8)
        cnt = C.c_size_t()
        _ffi.check(self._lib.pcp_tfrecord_scan(_ffi.ptr(self.buf), self.buf.size, 0, None, None, 0, C.byref(cnt)))
        self.offsets = np.empty(cnt.value, np.uint64)
        self.lengths = np.empty(cnt.value, np.uint64)
        if cnt.value:
            _ffi.check(self._lib.pcp_tfrecord_scan(_ffi.ptr(self.buf), self.buf.size, verify_crc, _ffi.ptr(self.offsets), _ffi.ptr(self.lengths),
                                                   cnt.value, C.byref(cnt)))

    def __len__(self):
        return int(self.offsets.size)

    def raw(self, i: int) -> bytes:
        o, l = int(self.offsets[i]), int(self.lengths[i])
        return self.buf[o:o + l].tobytes()

    def decode_into(self, i: int, width: int, obs: np.ndarray, part: np.ndarray, se3: np.ndarray) -> int:
        """Decode record i straight into caller-owned rows (obs (width,3) f32, part (width,) i32, se3 (3,3) f32)."""
        cl = C.c_int64()
        o, l = int(self.offsets[i]), int(self.lengths[i])
        _ffi.check(self._lib.pcp_example_decode(C.c_void_p(self.buf.ctypes.data + o), l, width, _ffi.ptr(obs), C.byref(cl), _ffi.ptr(part), _ffi.ptr(se3)))
        return int(cl.value)

    def record_ptr(self, i: int):
        """(address, length) of record i's payload inside the mapped file (for pcp_example_decode_batch)."""
        return self.buf.ctypes.data + int(self.offsets[i]), int(self.lengths[i])

    def decode(self, i: int, width: int):
        obs = np.empty((width, 3), np.float32)
        part = np.empty((width,), np.int32)
        se3 = np.empty((3, 3), np.float32)
        cl = self.decode_into(i, width, obs, part, se3)
        return obs, cl, part, se3
