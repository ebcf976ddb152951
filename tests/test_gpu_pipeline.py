"""GPU parity of BASELINE config 5 (raw scan -> voxel grid -> resize to width -> PointNet classify) against the oracle
chain voxel_downsample -> adjust_to_input_width -> forward, plus the host (pipelined H2D) path on a multi-chunk batch."""
import numpy as np
import pytest
import torch

from oracle import pointnet_numpy as onp
from oracle import pointnet_torch as opt
from oracle import sampling as osm
from tests import synth

pytestmark = pytest.mark.gpu


def rel_err(a, ref):
    return float(np.abs(a.astype(np.float64) - ref.astype(np.float64)).max() / max(np.abs(ref).max(), 1e-30))


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-3), ("bf16", 2e-2)])
def test_voxel_resize_classify_matches_oracle(lib, precision, tol):
    from pointcloudprocessing_b200 import pipeline
    from pointcloudprocessing_b200.pointnet import PointNet
    b, nraw, width, voxel = 3, 20000, 2048, 0.7
    scans = synth.lidar_scans(5005, b, nraw)
    p = onp.make_params(23, 12, seed=7)
    m = PointNet(23, 12, 0.3, 42, precision=precision)
    m.set_weights_dict(p)
    xd = torch.from_numpy(scans).cuda()
    got_x = pipeline.voxel_resize(xd, voxel, width).cpu().numpy()
    ref_x = np.empty((b, width, 3), np.float32)
    for c in range(b):
        cen, cnt, keys = osm.voxel_downsample(scans[c], voxel)
        assert cen.shape[0] > width                       # truncate branch (PointCloudSet.py:455-456)
        ref_x[c] = cen[:width]
    assert np.abs(got_x - ref_x).max() <= 1e-5 * np.abs(ref_x).max()
    cls, _, R = pipeline.classify_scans(m, xd, voxel, width)
    with torch.no_grad():
        ref = opt.forward(torch.from_numpy(ref_x), opt.to_torch(p), heads=("cls",))
    assert rel_err(cls.cpu().numpy(), ref["cls"].numpy()) <= tol
    assert rel_err(R.cpu().numpy(), ref["R"].numpy()) <= tol


def test_voxel_resize_pads_sparse_clouds(lib):
    from pointcloudprocessing_b200 import pipeline
    scans = synth.lidar_scans(77, 2, 3000)
    out = pipeline.voxel_resize(torch.from_numpy(scans).cuda(), 1.0, 4096, seed=3).cpu().numpy()
    for c in range(2):
        cen, _, _ = osm.voxel_downsample(scans[c], 1.0)
        v = cen.shape[0]
        assert v < 4096
        assert np.abs(out[c, :v] - cen).max() <= 1e-5 * np.abs(cen).max()
        # every padded row is a duplicate of one of the cloud's own voxels (PointCloudSet.py:459-467)
        pad = out[c, v:]
        d = np.abs(pad[:, None, :] - out[c, :v][None, :, :]).max(-1).min(1)
        assert d.max() == 0.0


def test_host_path_is_pipelined_over_chunks_and_exact(lib):
    """pcp_model_predict_host cuts a large batch into ~1 M-point chunks (copy stream || compute stream): results must be
    identical to the single-launch device path, cloud for cloud."""
    from pointcloudprocessing_b200.pointnet import PointNet
    p = onp.make_params(23, 12, seed=7)
    m = PointNet(23, 12, 0.3, 42, precision="bf16")
    m.set_weights_dict(p)
    x = synth.gaussian_clouds(9, 2500, 1024)               # 2.56 M points -> 3 chunks
    c_host, _, r_host = m(x, heads=("cls",))
    xd = torch.from_numpy(x).cuda()
    for lo, hi in ((0, 1024), (1024, 2048), (2048, 2500)):
        c_dev, _, r_dev = m(xd[lo:hi], heads=("cls",))
        assert np.array_equal(c_dev.cpu().numpy(), c_host[lo:hi]) and np.array_equal(r_dev.cpu().numpy(), r_host[lo:hi])
    c2, s2, _ = m(x[:40])                                   # with the segmentation head streamed back per chunk
    c3, s3, _ = m(xd[:40])
    assert np.array_equal(s3.cpu().numpy(), s2) and np.array_equal(c3.cpu().numpy(), c2)
