"""GPU parity: preprocessing kernels vs the oracle, called through the C-ABI (ops.py is a ctypes shim)."""
import numpy as np
import pytest
import torch

from oracle import pointnet_numpy as onp
from oracle import sampling as osm
from tests import synth

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _ulp_err(a, ref, scale):
    return np.abs(a.astype(np.float64) - ref.astype(np.float64)).max() / (np.finfo(np.float32).eps * scale)


# (2, 100000), (160, 16384), (3, 40000): the ticketed slice kernel (one round / more tickets than resident CTAs / uneven last slice)
@pytest.mark.parametrize("b,n", [(32, 1024), (8, 8192), (3, 1000), (5, 37), (2, 100000), (1, 1), (64, 2048), (160, 16384), (3, 40000), (2, 16386), (1, 300000)])
def test_normalize_matches_oracle(lib, b, n):
    from pointcloudprocessing_b200 import ops
    x = synth.gaussian_clouds(1001 + n, b, n)
    y, (c, s) = ops.normalize(_dev(x), return_stats=True)
    ry, (rc, rs) = onp.normalize(x)
    # centroid: double-accumulated on both sides -> identical up to the final rounding
    assert _ulp_err(c.cpu().numpy(), rc, np.abs(rc).max() + 1e-30) <= 1.0
    assert _ulp_err(s.cpu().numpy(), rs, np.abs(rs).max()) <= 2.0
    # outputs live in the unit ball; tolerance = a few ulp of the *input* magnitude relative to scale
    tol_ulps = 4.0 * (np.abs(x).max() / rs.min())
    assert _ulp_err(y.cpu().numpy(), ry, 1.0) <= max(tol_ulps, 4.0)
    if n > 1:
        assert np.array_equal(y.cpu().numpy(), ry) or _ulp_err(y.cpu().numpy(), ry, 1.0) <= max(tol_ulps, 4.0)


@pytest.mark.parametrize("b,n", [(2, 100000), (24, 16384), (4, 8192)])
def test_normalize_write_is_the_true_divide_of_its_own_statistics(lib, b, n):
    """PointNet.py:703-704: (x - centroid) / scale with IEEE subtract and divide — bit for bit given the kernel's own centroid
    and scale (the slice kernel divides through a corrected reciprocal; this pins it to the correctly rounded quotient)."""
    from pointcloudprocessing_b200 import ops
    x = synth.gaussian_clouds(99 + n, b, n)
    x[0, :64] = x[0, 0]                                       # exact zeros after centring are fine; tiny quotients too
    x[-1, 5] *= np.float32(1e-6)
    y, (c, s) = ops.normalize(_dev(x), return_stats=True)
    c, s = c.cpu().numpy().reshape(b, 1, 3), s.cpu().numpy().reshape(b, 1, 1)
    assert np.array_equal(y.cpu().numpy(), (x - c) / s)


def test_normalize_scratch_ownership(lib):
    """The three forms of the sliced normalisation — caller-owned workspace (what ops.normalize uses), the library's per-device ring
    with far more calls in flight on different streams than the ring has slots, and the static-array form — give the same bits."""
    from pointcloudprocessing_b200 import _ffi, ops
    import ctypes as C
    L = _ffi.load()
    b, n = 6, 40_000
    xs = [_dev(synth.gaussian_clouds(500 + i, b, n)) for i in range(20)]
    want = [ops.normalize(x) for x in xs]                      # pcp_normalize_ws_f32
    need = C.c_size_t(0)
    _ffi.check(L.pcp_normalize_workspace_bytes(b, n, C.byref(need)))
    assert need.value >= 64 + 48 * b
    _ffi.check(L.pcp_normalize_workspace_bytes(32, 1024, C.byref(need)))
    assert need.value == 0                                     # cluster-per-cloud path: no scratch
    streams = [torch.cuda.Stream() for _ in range(5)]
    got = [torch.empty_like(x) for x in xs]
    torch.cuda.synchronize()
    for i, x in enumerate(xs):                                 # 20 calls, 5 streams, 8 ring slots, nothing synchronises in between
        st = streams[i % len(streams)]
        _ffi.check(L.pcp_normalize_f32(_ffi.ptr(x), b, n, _ffi.ptr(got[i]), None, None, C.c_void_p(st.cuda_stream)))
    torch.cuda.synchronize()
    for g, w in zip(got, want):
        assert torch.equal(g, w)
    y = torch.empty_like(xs[0])
    _ffi.check(L.pcp_normalize_noalloc_f32(_ffi.ptr(xs[0]), b, n, _ffi.ptr(y), None, None, _ffi.stream_ptr()))
    assert torch.equal(y, want[0])
    small = torch.empty(16, device="cuda", dtype=torch.uint8)
    rc = L.pcp_normalize_ws_f32(_ffi.ptr(xs[0]), b, n, _ffi.ptr(y), None, None, _ffi.ptr(small), 16, _ffi.stream_ptr())
    assert rc != 0 and b"workspace" in L.pcp_last_error()


def test_normalize_degenerate(lib):
    from pointcloudprocessing_b200 import ops
    same = np.tile(np.array([[[3.0, -2.0, 5.0]]], np.float32), (2, 9, 1))
    y, (c, s) = ops.normalize(_dev(same), return_stats=True)
    assert np.array_equal(y.cpu().numpy(), np.zeros_like(same))
    assert np.all(s.cpu().numpy() == np.float32(1e-7))          # PointNet.py:701


def test_jitter_explicit_noise_bit_exact(lib):
    from pointcloudprocessing_b200 import ops
    rng = np.random.default_rng(3)
    for b, n in [(4, 1024), (3, 333)]:
        x = synth.gaussian_clouds(7, b, n)
        noise = rng.standard_normal((b, n, 3)).astype(np.float32)
        sd = np.array([0.1, 0.05, 0.2], np.float32)
        y = ops.jitter(_dev(x), sd, noise=_dev(noise)).cpu().numpy()
        assert np.array_equal(y, onp.jitter(x, noise, sd))       # PointCloudSet.py:335-337


def test_philox_noise_matches_oracle_bits_and_is_normal(lib):
    from pointcloudprocessing_b200 import ops
    npts, seed, off = 100_000, 0x1234_5678_9ABC, 77
    z = ops.philox_normal(npts, seed, off).cpu().numpy()
    bits = osm.jitter_noise_bits(npts, seed, off)
    u = lambda r, plus: ((r >> 8).astype(np.float32) + np.float32(plus)) * np.float32(1.0 / 16777216.0)
    ra = np.sqrt(-2.0 * np.log(u(bits[:, 0], 1).astype(np.float64)))
    rb = np.sqrt(-2.0 * np.log(u(bits[:, 2], 1).astype(np.float64)))
    ref = np.stack([ra * np.cos(2 * np.pi * u(bits[:, 1], 0).astype(np.float64)),
                    ra * np.sin(2 * np.pi * u(bits[:, 1], 0).astype(np.float64)),
                    rb * np.cos(2 * np.pi * u(bits[:, 3], 0).astype(np.float64))], 1)
    assert np.abs(z - ref).max() < 2e-5
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1.0) < 0.01
    # jitter(seed) == x + philox_normal(seed) * sd, bit for bit, and shard-invariant through `offset`
    x = synth.gaussian_clouds(9, 4, 25000)
    sd = np.array([0.1, 0.1, 0.1], np.float32)
    y = ops.jitter(_dev(x), sd, seed=seed, offset=off).cpu().numpy()
    assert np.array_equal(y, onp.jitter(x, z.reshape(4, 25000, 3), sd))
    y2 = ops.jitter(_dev(x[2:]), sd, seed=seed, offset=off + 2 * 25000).cpu().numpy()
    assert np.array_equal(y2, y[2:])


def test_rotate_bit_exact(lib):
    from pointcloudprocessing_b200 import ops
    rng = np.random.default_rng(11)
    for b, n in [(6, 2048), (2, 777)]:
        x = synth.gaussian_clouds(5, b, n)
        R = np.stack([onp.get_dcm(*rng.uniform(-180, 180, 3)) for _ in range(b)]).astype(np.float32)
        y = ops.rotate(_dev(x), _dev(R)).cpu().numpy()
        assert np.array_equal(y, onp.rotate(x, R))


def test_resize_to_width(lib):
    from pointcloudprocessing_b200 import ops
    x = synth.gaussian_clouds(2, 3, 500)
    lab = np.random.default_rng(0).integers(0, 12, (3, 500)).astype(np.int32)
    nv = np.array([500, 120, 313], np.int32)
    # truncate (PointCloudSet.py:455-456)
    y, l = ops.resize_to_width(_dev(x), torch.from_numpy(nv), 100, labels=_dev(lab))
    assert np.array_equal(y.cpu().numpy(), x[:, :100]) and np.array_equal(l.cpu().numpy(), lab[:, :100])
    # pad (PointCloudSet.py:459-467): first N kept, the rest are duplicates of existing points with their labels
    y, l = ops.resize_to_width(_dev(x), torch.from_numpy(nv), 500, seed=5, labels=_dev(lab))
    y, l = y.cpu().numpy(), l.cpu().numpy()
    for c in range(3):
        N = nv[c]
        assert np.array_equal(y[c, :N], x[c, :N]) and np.array_equal(l[c, :N], lab[c, :N])
        if N < 500:
            ctr = np.zeros((500 - N, 4), np.uint32)
            ctr[:, 0] = np.arange(N, 500); ctr[:, 2] = c; ctr[:, 3] = 0x5e512e
            r = osm.philox4x32_10(ctr, np.array([5, 0], np.uint32))[:, 0].astype(np.uint64)
            idx = ((r * np.uint64(N)) >> np.uint64(32)).astype(np.int64)
            o, ll = onp.adjust_to_input_width(x[c, :N], lab[c, :N], 500, pad_indices=idx)
            assert np.array_equal(y[c], o) and np.array_equal(l[c], ll)
            assert idx.min() >= 0 and idx.max() < N and len(np.unique(idx)) > min(N, 500 - N) // 4


# (75, 20000, 24) / (76, 20000, 1): more clusters than SMs -> two clouds per cluster (registers + shared memory), odd / even batch
@pytest.mark.parametrize("b,n,s", [(4, 1000, 64), (2, 20000, 128), (3, 4096, 256), (1, 7, 7), (2, 70000, 32), (75, 20000, 24), (76, 20000, 1)])
def test_fps_bit_exact(lib, b, n, s):
    from pointcloudprocessing_b200 import ops
    x = synth.lidar_scans(2002 + n, b, n, dup_frac=0.02)
    idx = ops.farthest_point_sample(_dev(x), s).cpu().numpy()
    assert np.array_equal(idx, osm.fps_batch(x, s))


def test_fps_ties_and_identical_points(lib):
    from pointcloudprocessing_b200 import ops
    pts = np.array([[[0, 0, 0], [2, 0, 0], [-2, 0, 0], [1, 0, 0], [-1, 0, 0]]], np.float32)
    assert ops.farthest_point_sample(_dev(pts), 5).cpu().numpy().tolist() == [[0, 1, 2, 3, 4]]
    ones = np.ones((2, 600, 3), np.float32)
    assert ops.farthest_point_sample(_dev(ones), 4, start=5).cpu().numpy().tolist() == [[5, 0, 0, 0]] * 2
    cube = np.array([[[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)]], np.float32)
    assert np.array_equal(ops.farthest_point_sample(_dev(cube), 8).cpu().numpy()[0], osm.fps(cube[0], 8))


def test_fps_full_size_c2(lib, oracle_c):
    """BASELINE config 2: 64 x 100 000 -> 1024.  Two clouds bit-exact vs the C oracle; all clouds through
    size-independent properties (valid, distinct indices; min-distance sequence non-increasing)."""
    import ctypes
    from pointcloudprocessing_b200 import ops
    b, n, s = 64, 100_000, 1024
    x = synth.lidar_scans(2002, b, n)
    idx = ops.farthest_point_sample(_dev(x), s).cpu().numpy()
    ref = np.empty((2, s), np.int32)
    oracle_c.oracle_fps_batch_f32(x[:2].ctypes.data_as(ctypes.c_void_p), 2, ctypes.c_int64(n), s, 0, ref.ctypes.data_as(ctypes.c_void_p))
    assert np.array_equal(idx[:2], ref)
    assert idx.min() >= 0 and idx.max() < n and (idx[:, 0] == 0).all()
    for c in (5, 63):
        sel = x[c][idx[c]].astype(np.float64)
        d = np.full(s, np.inf)
        mind = []
        for i in range(1, s):
            d = np.minimum(d, ((sel - sel[i - 1]) ** 2).sum(1))
            mind.append(((sel[:i] - sel[i]) ** 2).sum(1).min())
        assert all(mind[i] >= mind[i + 1] * (1 - 1e-6) for i in range(len(mind) - 1))


def test_gather_points(lib):
    from pointcloudprocessing_b200 import ops
    x = synth.gaussian_clouds(2, 3, 500)
    idx = np.random.default_rng(1).integers(0, 500, (3, 64)).astype(np.int32)
    out = ops.gather_points(_dev(x), _dev(idx)).cpu().numpy()
    assert np.array_equal(out, np.take_along_axis(x, idx[..., None].astype(np.int64), 1))


@pytest.mark.parametrize("b,n,voxel,origin", [(2, 4000, 0.75, None), (3, 20000, 0.5, None), (2, 5000, 1.0, (0.0, -20.0, -10.0)), (1, 131072, 0.35, None),
                                              (3, 4001, 0.75, None), (2, 130, 2.0, None),   # n % 4 != 0: scalar loads on the fast path; fewer than one warp run

                                              # bounding grid <= 32 n cells: bitmap-rank fast path; tiny voxels on few points: sort path
                                              (4, 100000, 0.7, None), (2, 50000, 2.0, (-3.0, 5.0, 1.0)), (2, 300, 0.05, None), (1, 1, 0.5, None)])
def test_voxel_matches_oracle(lib, b, n, voxel, origin):
    from pointcloudprocessing_b200 import ops
    x = synth.lidar_scans(5005 + n, b, n)
    cent, cnt, nv = ops.voxel_downsample(_dev(x), voxel, origin)
    cent, cnt, nv = cent.cpu().numpy(), cnt.cpu().numpy(), nv.cpu().numpy()
    for c in range(b):
        rc, rn, _ = osm.voxel_downsample(x[c], voxel, origin)
        v = rc.shape[0]
        assert nv[c] == v
        assert np.array_equal(cnt[c, :v], rn) and (cnt[c, v:] == 0).all()
        assert np.array_equal(cent[c, :v], rc)
    assert cnt.sum(axis=1).tolist() == [n] * b


def test_voxel_grid_caller_owned_capacity_bounded(lib):
    """pcp_voxel_compact_f32 through ops.VoxelGrid: asynchronous, caller-owned workspace and outputs, rows past
    min(nvox, capacity) untouched, buffers reusable across calls, truncation visible through nvox > capacity, -2 for a batch
    the bitmap path does not cover (then pcp_voxel_f32 takes it)."""
    from pointcloudprocessing_b200 import ops
    b, n, cap = 5, 30000, 6000
    vg = ops.VoxelGrid(b, n, cap)
    truncated = covered = 0
    for seed, voxel, origin in ((11, 0.7, None), (12, 0.9, (-2.0, -40.0, -12.0)), (13, 0.6, None), (14, 0.35, None)):
        x = synth.lidar_scans(seed, b, n)
        vg.centroids.fill_(-7.0); vg.counts.fill_(-7)
        cent, cnt, nv = vg(_dev(x), voxel, origin)
        cent, cnt, nv = cent.cpu().numpy(), cnt.cpu().numpy(), nv.cpu().numpy()
        if (nv == -2).any():                          # bounding grid above 32 n cells: the whole batch is handed back untouched
            assert (nv == -2).all() and (cent == -7.0).all() and (cnt == -7).all()
            continue
        covered += 1
        for c in range(b):
            rc, rn, _ = osm.voxel_downsample(x[c], voxel, origin)
            v = rc.shape[0]
            assert nv[c] == v
            k = min(v, cap)
            truncated += v > cap
            assert np.array_equal(cent[c, :k], rc[:k]) and np.array_equal(cnt[c, :k], rn[:k])
            assert (cent[c, k:] == -7.0).all() and (cnt[c, k:] == -7).all()
    assert covered >= 3 and truncated > 0             # the capacity really was exceeded somewhere
    # a grid outside the bitmap path (more than 32 n cells): 1 cm voxels on a 300-point cloud
    vg2 = ops.VoxelGrid(2, 300, 300)
    x = synth.lidar_scans(5, 2, 300)
    _, _, nv = vg2(_dev(x), 0.01)
    assert nv.cpu().numpy().tolist() == [-2, -2]
    cent, cnt, nv = ops.voxel_downsample(_dev(x), 0.01)
    rc, rn, _ = osm.voxel_downsample(x[1], 0.01)
    assert int(nv[1]) == rc.shape[0] and np.array_equal(cent[1, :rc.shape[0]].cpu().numpy(), rc)


def test_voxel_non_finite_point_flags_the_cloud(lib):
    """A NaN coordinate makes the cloud's voxel count -1 (oracle: ValueError) on either path; other clouds are unaffected."""
    from pointcloudprocessing_b200 import ops
    for n, voxel in ((60000, 0.7), (500, 0.05)):
        x = synth.lidar_scans(77, 2, n)
        x[1, n // 3, 1] = np.nan
        cent, cnt, nv = ops.voxel_downsample(_dev(x), voxel)
        nv = nv.cpu().numpy()
        rc, rn, _ = osm.voxel_downsample(x[0], voxel)
        assert nv[0] == rc.shape[0] and nv[1] == -1
        assert np.array_equal(cent[0, :nv[0]].cpu().numpy(), rc) and np.array_equal(cnt[0, :nv[0]].cpu().numpy(), rn)


def test_golden_sampling_fixture(lib):
    from pathlib import Path
    from pointcloudprocessing_b200 import ops
    s = np.load(Path(__file__).parent / "golden" / "sampling_small.npz")
    pts = synth.lidar_scans(int(s["seed"]), 2, int(s["n"]))
    assert np.array_equal(ops.farthest_point_sample(_dev(pts), int(s["s"])).cpu().numpy(), s["fps_idx"])
    cent, cnt, nv = ops.voxel_downsample(_dev(pts), float(s["voxel"]))
    v = int(nv[0])
    assert v == s["vox_cent"].shape[0] and np.array_equal(cent[0, :v].cpu().numpy(), s["vox_cent"])
