"""Voxel downsample of BASELINE config-2-sized scans (64 x 100 000 points, 0.7 m) for ncu launch lists / timing."""
import sys, time
import torch
sys.path.insert(0, ".")
from tests import synth
from pointcloudprocessing_b200 import ops

b, n = 64, 100_000
x = torch.from_numpy(synth.lidar_scans(2002, b, n)).cuda()
for i in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    cent, cnt, nv = ops.voxel_downsample(x, 0.7)
    torch.cuda.synchronize()
    print("iter", i, "%.3f ms" % ((time.perf_counter() - t0) * 1e3), "voxels/cloud", float(nv.float().mean()), flush=True)
