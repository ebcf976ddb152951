"""GPU: the PointCloudSet device iterator (tf.data semantics of pointcloud/PointCloudSet.py:341-391: [interleave] -> shuffle ->
repeat -> parse + jitter -> batch -> prefetch) — producer thread, pinned ring, copy stream, jitter on the device."""
import numpy as np
import pytest
import torch

from pointcloudprocessing_b200 import ops
from pointcloudprocessing_b200.pointcloud.PointCloudSet import PointCloudSet

pytestmark = pytest.mark.gpu

CLASSES = {"kc46": 0, "f15": 1}
PARTS = {"wing": 0, "fuselage": 1, "engine": 2}


def make(tmp_path, jitter, n=64, bs=8, count=40):
    pcs = PointCloudSet("dev_it", CLASSES, PARTS, n, jitter_stdev_m=np.array(jitter, np.float64), val=0.1, test=0.1, batch_size=bs, rand_seed=11,
                        print_func=lambda *a, **k: None, data_path=str(tmp_path) + "/")
    rng = np.random.default_rng(1)
    for s in range(3):                                   # three files per split: the train interleave has something to interleave
        obs = rng.normal(0, 20, (count, n, 3))
        obs[:, 0, 0] = 1000.0 * s + np.arange(count)     # a tag per sample
        pcs.add_data(f"set{s}", obs, rng.integers(0, 2, count), rng.integers(0, 3, (count, n)), np.tile(np.eye(3), (count, 1, 1)), shuffle_points=False)
    return pcs


def test_device_iterator_equals_host_iterator(lib, tmp_path):
    pcs = make(tmp_path, (0.0, 0.0, 0.0))
    host = pcs.get_train_set()
    dev = pcs.get_train_set(device="cuda")
    kept = []
    for _ in range(30):                                  # > 2 passes over 96 train samples, > ring depth
        xh, yh = next(host)
        xd, yd = next(dev)
        assert np.array_equal(xd.cpu().numpy(), xh)
        for k in yh:
            assert np.array_equal(yd[k].cpu().numpy(), yh[k])
        kept.append((xd, xh))
    # batches handed out earlier are the caller's own: the ring was recycled many times since
    for xd, xh in kept:
        assert np.array_equal(xd.cpu().numpy(), xh)
    # every pass sees every train sample exactly once
    n_train = pcs._data_size['train']['count']
    tags = np.concatenate([xh[:, 0, 0] for _, xh in kept])[:2 * n_train]
    assert len(set(tags[:n_train].tolist())) == n_train and sorted(tags[:n_train]) == sorted(tags[n_train:])
    dev.close()


def test_device_iterator_jitter_is_seeded_philox_and_differs_between_splits(lib, tmp_path):
    sd = (0.1, 0.2, 0.05)
    pcs = make(tmp_path, sd)
    clean = pcs.get_val_set(jitter=False)
    a, b = pcs.get_val_set(device="cuda"), pcs.get_val_set(device="cuda")
    bs, w = 8, 64
    for step in range(3):
        x0, _ = next(clean)
        xa, _ = next(a)
        xb, _ = next(b)
        assert torch.equal(xa, xb)                       # same set seed, same split: the same noise
        # y = x + N(0,1) * stdev (separate roundings) with N from Philox4x32-10 keyed by (set seed + split, global point index):
        # pcp_philox_normal_f32 is pinned to the Random123 vectors / Box-Muller in tests/test_gpu_preprocess.py
        noise = ops.philox_normal(bs * w, pcs._random_seed + 1 + 1_000_003, offset=step * bs * w).cpu().numpy().reshape(bs, w, 3)
        ref = (x0 + (noise * np.asarray(sd, np.float32)).astype(np.float32)).astype(np.float32)
        assert np.array_equal(xa.cpu().numpy(), ref)
    xt, _ = next(pcs.get_test_set(device="cuda"))
    xv, _ = next(pcs.get_val_set(device="cuda"))
    assert xt.shape == xv.shape
    a.close(); b.close()
