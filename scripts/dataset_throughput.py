"""Development aid / profiles: records per second of the PointCloudSet device iterator (decode on host threads -> pinned ring ->
H2D on a copy stream -> jitter on the device) at BASELINE config 3's shape (batch 256 x 1024 points), alone and while a
training step consumes it."""
import sys
import tempfile
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from pointcloudprocessing_b200.pointcloud.PointCloudSet import PointCloudSet  # noqa: E402

classes = {f"c{i}": i for i in range(23)}
parts = {f"p{i}": i for i in range(12)}
n, bs, count = 1024, 256, 4096
with tempfile.TemporaryDirectory() as d:
    pcs = PointCloudSet("thr", classes, parts, n, jitter_stdev_m=np.array([0.1, 0.1, 0.1]), val=0.1, test=0.05, batch_size=bs, rand_seed=5,
                        print_func=lambda *a, **k: None, data_path=d + "/")
    rng = np.random.default_rng(0)
    for s in range(2):
        pcs.add_data(f"set{s}", rng.normal(0, 30, (count // 2, n, 3)), rng.integers(0, 23, count // 2), rng.integers(0, 12, (count // 2, n)),
                     np.tile(np.eye(3), (count // 2, 1, 1)), shuffle_points=False)
    for threads in (1, 4, 8):
        it = pcs.get_train_set(device="cuda", decode_threads=threads)
        for _ in range(3):
            next(it)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        steps = 40
        for _ in range(steps):
            x, y = next(it)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        print(f"device iterator, {threads} decode threads: {steps * bs / dt:9.0f} clouds/s  ({dt / steps * 1e3:.2f} ms per batch of {bs} x {n})")
        it.close()
    # host iterator (no device): decode only
    it = pcs.get_train_set(jitter=False, decode_threads=8)
    next(it)
    t0 = time.perf_counter()
    for _ in range(20):
        next(it)
    dt = time.perf_counter() - t0
    print(f"host iterator, 8 decode threads: {20 * bs / dt:9.0f} clouds/s")
