"""Writes the golden fixtures under tests/golden/ from the oracle (run from the repo root:
`python tests/golden/make_golden.py`).  The reference itself cannot be imported here (TensorFlow /
Keras absent), so these are oracle outputs — "parity unpinned" still applies; the
fixtures pin the oracle and give the GPU box inputs/outputs that do not need /root/reference."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import pointnet_numpy as onp  # noqa: E402
from oracle import sampling as osm  # noqa: E402
from tests import synth  # noqa: E402

here = Path(__file__).parent
C, P, b, n, wseed, iseed = 23, 12, 2, 128, 7, 1001
p = onp.make_params(C, P, seed=wseed)
x = synth.gaussian_clouds(iseed, b, n)
o = onp.forward(x, p, return_intermediates=True)
np.savez_compressed(here / "pointnet_small.npz", C=C, P=P, b=b, n=n, weight_seed=wseed, input_seed=iseed,
                    **{k: o[k] for k in ("cls", "seg", "R", "cls_logits", "seg_logits", "global_features")})
seed, n2, s, voxel = 2002, 4000, 96, 0.75
pts = synth.lidar_scans(seed, 2, n2)
cent, cnt, keys = osm.voxel_downsample(pts[0], voxel)
np.savez_compressed(here / "sampling_small.npz", seed=seed, n=n2, s=s, voxel=voxel, fps_idx=osm.fps_batch(pts, s),
                    vox_cent=cent, vox_cnt=cnt, vox_keys=keys)
print("golden fixtures written")
