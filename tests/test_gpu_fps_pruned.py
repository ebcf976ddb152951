"""The pruned farthest-point sampler (pcp_fps_ws_f32: Morton-ordered cloud, bounding-sphere skipping) must return exactly the samples of
the unpruned kernel and of the plain-C oracle — including on clouds made of exact ties (duplicates, lattices, identical points), where
the tie-break by lowest ORIGINAL index has to survive the reordering."""
import ctypes

import numpy as np
import pytest
import torch

from tests import synth

pytestmark = pytest.mark.gpu


def plain(x, s, start):
    from pointcloudprocessing_b200 import _ffi
    b, n, _ = x.shape
    idx = torch.empty((b, s), device=x.device, dtype=torch.int32)
    _ffi.check(_ffi.load().pcp_fps_f32(_ffi.ptr(x), b, n, s, start, _ffi.ptr(idx), _ffi.stream_ptr()))
    return idx


def oracle(oracle_c, x, s, start):
    b, n, _ = x.shape
    ref = np.empty((b, s), np.int32)
    oracle_c.oracle_fps_batch_f32(x.ctypes.data_as(ctypes.c_void_p), b, ctypes.c_int64(n), s, start, ref.ctypes.data_as(ctypes.c_void_p))
    return ref


def make(kind, b, n):
    rng = np.random.default_rng(b * 1000 + n)
    if kind == "lidar":
        return synth.lidar_scans(77, b, n)
    if kind == "dup":
        x = synth.lidar_scans(78, b, n)
        x[:, n // 2:2 * (n // 2)] = x[:, :n // 2]
        return x
    if kind == "lattice":
        return np.round(rng.uniform(-4, 4, (b, n, 3))).astype(np.float32)
    if kind == "same":
        return np.full((b, n, 3), 1.5, np.float32)
    if kind == "offset":
        return (rng.normal(0, 1, (b, n, 3)) * 1e-3 + 1000.0).astype(np.float32)
    if kind == "line":                                   # degenerate bounding box: two axes have no extent
        x = np.zeros((b, n, 3), np.float32)
        x[..., 0] = rng.uniform(0, 100, (b, n))
        return x
    raise ValueError(kind)


@pytest.mark.parametrize("kind,b,n,s,start", [
    ("lidar", 40, 100_000, 256, 0), ("lidar", 64, 70_001, 128, 77), ("lidar", 150, 13_000, 128, 5), ("dup", 40, 90_001, 200, 3),
    ("lattice", 40, 60_000, 300, 0), ("same", 40, 52_000, 40, 5), ("offset", 40, 100_000, 128, 0), ("line", 40, 75_000, 128, 1),
    ("dup", 40, 40_000, 64, 0), ("lidar", 10, 120_000, 64, 2),          # clusters of 2 and of 8 blocks (the others: 1 and 4)
])
def test_pruned_fps_matches_plain_kernel_and_oracle(lib, oracle_c, kind, b, n, s, start):
    from pointcloudprocessing_b200 import ops
    x = make(kind, b, n)
    xd = torch.from_numpy(x).cuda()
    assert ops.fps_workspace_bytes(b, n, s) > 0, "case must take the pruned path"
    got = ops.farthest_point_sample(xd, s, start)
    assert torch.equal(got, plain(xd, s, start))
    k = min(b, 6)                                        # the scalar oracle on a few clouds (the plain kernel is pinned on all of them elsewhere)
    assert np.array_equal(got[:k].cpu().numpy(), oracle(oracle_c, x[:k], s, start))


def test_pruned_fps_caller_workspace_and_errors(lib):
    from pointcloudprocessing_b200 import _ffi, ops
    b, n, s = 40, 60_000, 64
    xd = torch.from_numpy(synth.lidar_scans(5, b, n)).cuda()
    need = ops.fps_workspace_bytes(b, n, s)
    ws = torch.empty(need, device="cuda", dtype=torch.uint8)
    a = ops.farthest_point_sample(xd, s, 0, workspace=ws)
    assert torch.equal(a, ops.farthest_point_sample(xd, s, 0, workspace=ws))        # scratch re-used as is
    assert torch.equal(a, plain(xd, s, 0))
    with pytest.raises(ValueError):
        ops.farthest_point_sample(xd, s, 0, workspace=ws[: need // 2])
    idx = torch.empty((b, s), device="cuda", dtype=torch.int32)
    rc = _ffi.load().pcp_fps_ws_f32(_ffi.ptr(xd), b, n, s, 0, _ffi.ptr(idx), None, 0, _ffi.stream_ptr())
    assert rc != 0 and b"workspace" in _ffi.load().pcp_last_error()
    # few / small clouds: no workspace wanted, the plain kernel runs
    assert ops.fps_workspace_bytes(2, 4096, 64) == 0
    small = torch.from_numpy(synth.gaussian_clouds(3, 2, 4096)).cuda()
    assert torch.equal(ops.farthest_point_sample(small, 64), plain(small, 64, 0))
