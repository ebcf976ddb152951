import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from oracle import pointnet_numpy as onp
from pointcloudprocessing_b200.pointnet import PointNet
from tests import synth
m = PointNet(23, 12, 0.3, 42, precision="bf16"); m.set_weights_dict(onp.make_params(23, 12))
x = torch.from_numpy(synth.gaussian_clouds(1, 128, 2048)).cuda()
for _ in range(3):
    m(x)
torch.cuda.synchronize()
