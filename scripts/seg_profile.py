"""BASELINE config 4 (segmentation inference, 128 x 2048 points, all outputs) for ncu launch lists / timing."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from oracle import pointnet_numpy as onp          # parameter init only
from tests import synth
from pointcloudprocessing_b200.pointnet import PointNet

b, n = 128, 2048
m = PointNet(23, 12, 0.3, 42, precision="bf16")
m.set_weights_dict(onp.make_params(23, 12, seed=7))
x = torch.from_numpy(synth.gaussian_clouds(4004, b, n)).cuda()
for i in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = m(x)
    torch.cuda.synchronize()
    print("iter", i, "%.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
