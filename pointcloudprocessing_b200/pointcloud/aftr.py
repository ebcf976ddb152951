"""AftrBurner frame text ("(x, y, z) class part" per line) outside PointCloudSet: the writer of the reference's toolkit,
Provizio.to_aftr_frame (point_cloud_toolkit/utils/Provizio.py:126-145), and a stand-alone parser, both over the native
host routines of libpcp_b200.so (csrc/host_io.cu; no GPU needed).  The reference's pure-Python loops manage ~47 frames/s on
ingest (examples/pointnet_train.ipynb cell 6)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .. import _ffi


def format_aftr_frame(points: np.ndarray, labels: np.ndarray = np.array([])) -> bytes:
    """The frame text to_aftr_frame writes, as bytes.  points (N, 3) float32 / float64 (other dtypes are converted to
    float64 like Python's float(); a float32 prints with the digits of its double value, as in the reference); labels: ()
    for none, (N,) or (N, k) of anything str() can print."""
    pts = np.asarray(points)
    if pts.dtype != np.float32:
        pts = pts.astype(np.float64)
    pts = np.ascontiguousarray(pts)
    n = pts.shape[0]
    ids, table, k = None, None, 0
    if labels.shape[0] > 0:
        lab = np.asarray(labels).reshape(n, -1)
        k = lab.shape[1]
        uniq, inv = np.unique(lab.astype(str), return_inverse=True)
        ids = np.ascontiguousarray(inv.reshape(n, k).astype(np.int32))
        table = "\n".join(uniq.tolist()).encode()
    lib = _ffi.load()
    need = C.c_size_t(0)
    lib.pcp_format_aftr_frame(_ffi.ptr(pts), int(pts.dtype == np.float64), n, _ffi.ptr(ids), k, table, None, 0, C.byref(need))
    buf = np.empty(max(1, need.value), np.uint8)
    _ffi.check(lib.pcp_format_aftr_frame(_ffi.ptr(pts), int(pts.dtype == np.float64), n, _ffi.ptr(ids), k, table, _ffi.ptr(buf), buf.size, C.byref(need)))
    return buf[:need.value].tobytes()


def to_aftr_frame(path: str, points: np.ndarray, labels: np.ndarray = np.array([]), print_func=print) -> None:
    """Provizio.to_aftr_frame: same checks, same messages, same file contents."""
    if len(points.shape) != 2 or points.shape[1] != 3:
        print_func(f"Unable to create aftr frame -> points vector must be shape (N, 3), not {points.shape}.")
        return
    if points.shape[0] != labels.shape[0] and labels.shape[0] != 0:
        print_func(f"Unable to create aftr frame -> if labels are available, the number of labels much match the number of points. "
                   f"Currently there are {points.shape[0]} points and {labels.shape[0]} labels.")
        return
    if os.path.isdir(os.path.dirname(path)):
        with open(path, 'wb') as f:
            f.write(format_aftr_frame(points, labels))
    else:
        print_func("Unable to create aftr frame -> path does not exist.")


def parse_aftr_frame(text: bytes, class_labels, part_labels):
    """One frame text -> (xyz (N, 3) float64, class index, part indices (N,) int32, number of skipped non-finite points) with the
    semantics of PointCloudSet.add_from_aftr_output (pointcloud/PointCloudSet.py:161-198)."""
    cap = max(1, text.count(b'\n') + 1)
    xyz = np.empty((cap, 3), np.float64)
    part = np.empty((cap,), np.int32)
    cnt, bad, cls = C.c_size_t(), C.c_size_t(), C.c_int32()
    lib = _ffi.load()
    _ffi.check(lib.pcp_parse_aftr_frame(text, len(text), "\n".join(class_labels).encode(), "\n".join(part_labels).encode(), _ffi.ptr(xyz),
                                        _ffi.ptr(part), cap, C.byref(cnt), C.byref(cls), C.byref(bad)))
    return xyz[:cnt.value], int(cls.value), part[:cnt.value], int(bad.value)
