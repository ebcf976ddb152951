"""BASELINE config 1 batch (32 x 1024, cls + R) latency, for ncu launch lists / timing."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from oracle import pointnet_numpy as onp          # parameter init only
from tests import synth
from pointcloudprocessing_b200.pointnet import PointNet

b, n = 32, 1024
m = PointNet(23, 12, 0.3, 42, precision="bf16")
m.set_weights_dict(onp.make_params(23, 12, seed=7))
x = torch.from_numpy(synth.gaussian_clouds(4004, b, n)).cuda()
for heads in (("cls",), ("cls", "seg")):
    for i in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(20):
            out = m(x, heads=heads)
        torch.cuda.synchronize()
        print(heads, "iter", i, "%.1f us" % ((time.perf_counter() - t0) / 20 * 1e6), flush=True)
