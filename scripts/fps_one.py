"""One pruned-FPS call on the bench shape (64 x 100 000 -> 1024), for ncu: python scripts/fps_one.py [b n s]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloudprocessing_b200 import ops  # noqa: E402
from tests import synth  # noqa: E402

b, n, s = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (64, 100_000, 1024)
x = torch.from_numpy(synth.lidar_scans(2002, b, n)).cuda()
for _ in range(2):
    idx = ops.farthest_point_sample(x, s)
torch.cuda.synchronize()
print(int(idx.sum()))
