"""Time the fp32_tc (1e-3 contract on tensor cores) classification forward at 1024 x 1024, like bench.py's fp32_contract leg."""
import sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from oracle import pointnet_numpy as onp          # parameter init only
from tests import synth
from pointcloudprocessing_b200.pointnet import PointNet

m = PointNet(40, 50, 0.3, 42, precision="fp32_tc")
m.set_weights_dict(onp.make_params(40, 50, seed=7))
x = torch.from_numpy(synth.gaussian_clouds(5005, 1024, 1024)).cuda()
for i in range(3):
    m(x, heads=("cls",))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(10):
    m(x, heads=("cls",))
e1.record(); torch.cuda.synchronize()
print("fp32_tc 1024 x 1024: %.3f ms" % (e0.elapsed_time(e1) / 10))
