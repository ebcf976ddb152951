"""One call each of normalize / voxel / FPS at 64 x 100 000 (after a warm-up call) for `ncu --set full` captures."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from pointcloudprocessing_b200 import ops
from tests import synth
x = torch.from_numpy(synth.lidar_scans(2002, 64, 100_000)).cuda()
for _ in range(2):
    ops.normalize(x); ops.voxel_downsample(x, 0.7); ops.farthest_point_sample(x, 1024)
torch.cuda.synchronize()
