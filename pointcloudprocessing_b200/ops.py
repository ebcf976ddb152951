"""Host-side operators over the C-ABI (include/pcp.h) — the preprocessing half of the hot path.

Inputs are CUDA torch tensors (device memory plumbing only; every FLOP runs in libpcp_b200.so).
Names follow the reference: normalisation = PointCloudNormalization (pointnet/PointNet.py:681-706),
jitter = PointCloudSet._parse_function (pointcloud/PointCloudSet.py:335-337), resize =
PointCloudSet._adjust_to_input_width (:443-470); FPS / voxel / rotation are the north_star operators.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _ffi


def _chk_cloud(x):
    if not (isinstance(x, torch.Tensor) and x.is_cuda and x.dtype == torch.float32 and x.dim() == 3 and x.shape[-1] == 3):
        raise TypeError("expected a CUDA float32 tensor of shape (b, n, 3)")
    return x.contiguous()


def normalize(x: torch.Tensor, return_stats: bool = False):
    """(x - centroid) / max(maxdist, 1e-7) per cloud; optionally (centroid (b,1,3), scale (b,1,1))."""
    x = _chk_cloud(x)
    b, n, _ = x.shape
    y = torch.empty_like(x)
    cen = torch.empty((b, 1, 3), device=x.device, dtype=torch.float32) if return_stats else None
    sc = torch.empty((b, 1, 1), device=x.device, dtype=torch.float32) if return_stats else None
    if b and n:
        lib = _ffi.load()
        need = C.c_size_t(0)
        _ffi.check(lib.pcp_normalize_workspace_bytes(b, n, C.byref(need)))
        # caller-owned scratch from torch's caching allocator: no cudaMalloc and no library-side ring in the steady state
        ws = torch.empty(need.value, device=x.device, dtype=torch.uint8) if need.value else None
        _ffi.check(lib.pcp_normalize_ws_f32(_ffi.ptr(x), b, n, _ffi.ptr(y), _ffi.ptr(cen), _ffi.ptr(sc), _ffi.ptr(ws), need.value, _ffi.stream_ptr()))
    return (y, (cen, sc)) if return_stats else y


def philox_normal(npoints: int, seed: int, offset: int = 0, device="cuda") -> torch.Tensor:
    out = torch.empty((npoints, 3), device=device, dtype=torch.float32)
    _ffi.check(_ffi.load().pcp_philox_normal_f32(npoints, seed, offset, _ffi.ptr(out), _ffi.stream_ptr()))
    return out


def jitter(x: torch.Tensor, stdev_m, seed: int = 0, offset: int = 0, noise: torch.Tensor | None = None) -> torch.Tensor:
    """x + N(0,1) * stdev_axis.  noise=None draws Philox noise indexed by global point id + offset."""
    x = _chk_cloud(x)
    sd = np.ascontiguousarray(np.asarray(stdev_m, np.float32).reshape(3))
    y = torch.empty_like(x)
    if noise is not None:
        noise = noise.contiguous()
        assert noise.shape == x.shape and noise.dtype == torch.float32 and noise.is_cuda
    _ffi.check(_ffi.load().pcp_jitter_f32(_ffi.ptr(x), x.shape[0] * x.shape[1], _ffi.fptr(sd), seed, offset,
                                          _ffi.ptr(noise), _ffi.ptr(y), _ffi.stream_ptr()))
    return y


def rotate(x: torch.Tensor, R: torch.Tensor) -> torch.Tensor:
    """y = x . R^T per cloud; R (b,3,3)."""
    x = _chk_cloud(x)
    R = R.to(device=x.device, dtype=torch.float32).contiguous()
    assert R.shape == (x.shape[0], 3, 3)
    y = torch.empty_like(x)
    _ffi.check(_ffi.load().pcp_rotate_f32(_ffi.ptr(x), x.shape[0], x.shape[1], _ffi.ptr(R), _ffi.ptr(y), _ffi.stream_ptr()))
    return y


def farthest_point_sample(x: torch.Tensor, s: int, start: int = 0, workspace: torch.Tensor | None = None) -> torch.Tensor:
    """(b, n, 3) -> (b, s) int32 indices, bit-exact with oracle/sampling.py:fps.  `workspace`: optional caller-owned uint8 scratch of
    fps_workspace_bytes(b, n, s) bytes for the pruned kernel (allocated here when omitted)."""
    x = _chk_cloud(x)
    b, n, _ = x.shape
    idx = torch.empty((b, s), device=x.device, dtype=torch.int32)
    lib = _ffi.load()
    need = C.c_size_t(0)
    _ffi.check(lib.pcp_fps_workspace_bytes(b, n, s, C.byref(need)))
    if workspace is None and need.value:
        workspace = torch.empty(need.value, device=x.device, dtype=torch.uint8)   # caching allocator: no cudaMalloc in the steady state
    if need.value and (workspace.numel() * workspace.element_size() < need.value or workspace.device != x.device):
        raise ValueError(f"farthest_point_sample: workspace of {need.value} bytes on {x.device} needed")
    _ffi.check(lib.pcp_fps_ws_f32(_ffi.ptr(x), b, n, s, start, _ffi.ptr(idx), _ffi.ptr(workspace) if need.value else None,
                                  need.value, _ffi.stream_ptr()))
    return idx


def fps_workspace_bytes(b: int, n: int, s: int) -> int:
    need = C.c_size_t(0)
    _ffi.check(_ffi.load().pcp_fps_workspace_bytes(b, n, s, C.byref(need)))
    return need.value


def gather_points(x: torch.Tensor, idx: torch.Tensor, labels: torch.Tensor | None = None):
    x = _chk_cloud(x)
    b, n, _ = x.shape
    idx = idx.contiguous()
    s = idx.shape[1]
    out = torch.empty((b, s, 3), device=x.device, dtype=torch.float32)
    lo = None
    if labels is not None:
        labels = labels.to(torch.int32).contiguous()
        lo = torch.empty((b, s), device=x.device, dtype=torch.int32)
    _ffi.check(_ffi.load().pcp_gather_points_f32(_ffi.ptr(x), _ffi.ptr(labels), b, n, _ffi.ptr(idx), s, _ffi.ptr(out),
                                                 _ffi.ptr(lo), _ffi.stream_ptr()))
    return (out, lo) if labels is not None else out


def resize_to_width(x: torch.Tensor, n_valid: torch.Tensor, width: int, seed: int = 0, labels: torch.Tensor | None = None):
    """_adjust_to_input_width on the device: x (b, n_max, 3) with n_valid[c] real points per cloud."""
    x = _chk_cloud(x)
    b, nmax, _ = x.shape
    nv = n_valid.to(device=x.device, dtype=torch.int32).contiguous()
    y = torch.empty((b, width, 3), device=x.device, dtype=torch.float32)
    lo = None
    if labels is not None:
        labels = labels.to(torch.int32).contiguous()
        lo = torch.empty((b, width), device=x.device, dtype=torch.int32)
    _ffi.check(_ffi.load().pcp_resize_to_width_f32(_ffi.ptr(x), _ffi.ptr(labels), b, nmax, _ffi.ptr(nv), width, seed,
                                                   _ffi.ptr(y), _ffi.ptr(lo), _ffi.stream_ptr()))
    return (y, lo) if labels is not None else y


class VoxelGrid:
    """Voxel-grid downsampler with caller-owned buffers (workspace, capacity-bounded outputs): every call is asynchronous on
    the current stream, allocates nothing and reads nothing back (pcp_voxel_compact_f32).  `nvox` reports per cloud the
    voxel count, -1 (cloud rejected: non-finite coordinate) or -2 (batch not covered by the bitmap path: use
    voxel_downsample, which falls back to the general path)."""

    def __init__(self, b: int, n: int, capacity: int | None = None, device="cuda", zero_fill: bool = False):
        self.b, self.n, self.capacity = int(b), int(n), int(capacity if capacity is not None else n)
        lib = _ffi.load()
        nbytes = C.c_size_t(0)
        _ffi.check(lib.pcp_voxel_workspace_bytes(self.b, self.n, C.byref(nbytes)))
        self._ws_bytes = nbytes.value
        self._ws = torch.empty((self._ws_bytes,), device=device, dtype=torch.uint8)
        alloc = torch.zeros if zero_fill else torch.empty
        self.centroids = alloc((self.b, self.capacity, 3), device=device, dtype=torch.float32)
        self.counts = alloc((self.b, self.capacity), device=device, dtype=torch.int32)
        self.nvox = torch.empty((self.b,), device=device, dtype=torch.int32)

    def __call__(self, x: torch.Tensor, voxel: float, origin=None):
        x = _chk_cloud(x)
        assert tuple(x.shape[:2]) == (self.b, self.n), (x.shape, self.b, self.n)
        org = None if origin is None else np.ascontiguousarray(np.asarray(origin, np.float32).reshape(3))
        _ffi.check(_ffi.load().pcp_voxel_compact_f32(_ffi.ptr(x), self.b, self.n, _ffi.fptr(org), float(voxel), self.capacity, _ffi.ptr(self.centroids),
                                                    _ffi.ptr(self.counts), _ffi.ptr(self.nvox), _ffi.ptr(self._ws), self._ws_bytes, _ffi.stream_ptr()))
        return self.centroids, self.counts, self.nvox


def voxel_downsample(x: torch.Tensor, voxel: float, origin=None, capacity: int | None = None):
    """Voxel-grid downsample: returns (centroids (b, capacity, 3), counts (b, capacity) int32, nvox (b,) int32); rows past
    nvox are zero; capacity defaults to n (a cloud cannot have more voxels than points).  Runs the bitmap path
    (pcp_voxel_compact_f32) and, for grids it does not cover, the general sort path (pcp_voxel_f32) — that decision reads
    nvox back, so this convenience form synchronises; VoxelGrid is the asynchronous form."""
    x = _chk_cloud(x)
    b, n, _ = x.shape
    cap = int(capacity if capacity is not None else n)
    if b == 0 or n == 0:
        return (torch.zeros((b, cap, 3), device=x.device), torch.zeros((b, cap), device=x.device, dtype=torch.int32),
                torch.zeros((b,), device=x.device, dtype=torch.int32))
    vg = VoxelGrid(b, n, cap, device=x.device, zero_fill=True)
    out, cnt, nv = vg(x, voxel, origin)
    if bool((nv == -2).any()):
        return _voxel_general(x, voxel, origin, cap)
    return out, cnt, nv


def _voxel_general(x: torch.Tensor, voxel: float, origin, cap: int):
    """pcp_voxel_f32: bitmap-rank path / radix-sort path, outputs padded to n rows per cloud."""
    b, n, _ = x.shape
    lib = _ffi.load()
    nbytes = C.c_size_t(0)
    _ffi.check(lib.pcp_voxel_workspace_bytes(b, n, C.byref(nbytes)))
    ws = torch.empty((nbytes.value,), device=x.device, dtype=torch.uint8)
    out = torch.empty((b, n, 3), device=x.device, dtype=torch.float32)      # every row is written (zero padded past nvox)
    cnt = torch.empty((b, n), device=x.device, dtype=torch.int32)
    nv = torch.empty((b,), device=x.device, dtype=torch.int32)
    org = None if origin is None else np.ascontiguousarray(np.asarray(origin, np.float32).reshape(3))
    _ffi.check(lib.pcp_voxel_f32(_ffi.ptr(x), b, n, _ffi.fptr(org), float(voxel), _ffi.ptr(out), _ffi.ptr(cnt), _ffi.ptr(nv),
                                 _ffi.ptr(ws), nbytes.value, _ffi.stream_ptr()))
    if cap < n:
        out, cnt = out[:, :cap].contiguous(), cnt[:, :cap].contiguous()
    elif cap > n:
        out = torch.nn.functional.pad(out, (0, 0, 0, cap - n))
        cnt = torch.nn.functional.pad(cnt, (0, cap - n))
    return out, cnt, nv


# ---- pose post-processing (PCA/utils/mat_ops.py, PCT/utils/calibration.py), batched on the device ----
def _f32_cuda(a, shape_tail):
    t = a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a, np.float32))
    t = t.to(device="cuda", dtype=torch.float32).contiguous()
    assert tuple(t.shape[1:]) == tuple(shape_tail), t.shape
    return t


def get_dcm(rpy_deg) -> torch.Tensor:
    """mat_ops.get_dcm for a batch: (b, 3) [roll, pitch, yaw] degrees -> (b, 3, 3)."""
    r = _f32_cuda(rpy_deg, (3,))
    out = torch.empty((r.shape[0], 3, 3), device=r.device, dtype=torch.float32)
    _ffi.check(_ffi.load().pcp_dcm_from_rpy_f32(_ffi.ptr(r), r.shape[0], _ffi.ptr(out), _ffi.stream_ptr()))
    return out


def get_roll_pitch_yaw_deg(dcm) -> torch.Tensor:
    """mat_ops.get_roll_pitch_yaw_deg for a batch: (b, 3, 3) -> (b, 3) [roll, pitch, yaw] degrees."""
    R = _f32_cuda(dcm, (3, 3))
    out = torch.empty((R.shape[0], 3), device=R.device, dtype=torch.float32)
    _ffi.check(_ffi.load().pcp_rpy_from_dcm_f32(_ffi.ptr(R), R.shape[0], _ffi.ptr(out), _ffi.stream_ptr()))
    return out


def reorthogonalize(dcm) -> torch.Tensor:
    """mat_ops.reorthogonalize for a batch (e.g. the model's predicted `se3`): nearest orthogonal matrix U V^T."""
    R = _f32_cuda(dcm, (3, 3))
    out = torch.empty_like(R)
    _ffi.check(_ffi.load().pcp_reorthogonalize_f32(_ffi.ptr(R), R.shape[0], _ffi.ptr(out), _ffi.stream_ptr()))
    return out


def solve_kabsch(truth_vecs, measured_vecs):
    """calibration.solve_kabsch for a batch of point-set pairs (b, n, 3): returns R (b,3,3), t (b,3), rmse (b,) with
    aligned = R @ measured + t."""
    T = truth_vecs if isinstance(truth_vecs, torch.Tensor) else torch.as_tensor(np.asarray(truth_vecs, np.float32))
    M = measured_vecs if isinstance(measured_vecs, torch.Tensor) else torch.as_tensor(np.asarray(measured_vecs, np.float32))
    T = T.to(device="cuda", dtype=torch.float32).contiguous()
    M = M.to(device="cuda", dtype=torch.float32).contiguous()
    assert T.dim() == 3 and T.shape == M.shape and T.shape[2] == 3
    b, n, _ = T.shape
    R = torch.empty((b, 3, 3), device=T.device, dtype=torch.float32)
    t = torch.empty((b, 3), device=T.device, dtype=torch.float32)
    rmse = torch.empty((b,), device=T.device, dtype=torch.float32)
    _ffi.check(_ffi.load().pcp_kabsch_f32(_ffi.ptr(T), _ffi.ptr(M), b, n, _ffi.ptr(R), _ffi.ptr(t), _ffi.ptr(rmse), _ffi.stream_ptr()))
    return R, t, rmse
