"""Small-case run of the round-2 kernels (pruned FPS at cluster sizes 1 and 2 with forced pruning and ties, fused dense tails, staged host path,\nworkspace normalise) against their plain counterparts; meant for compute-sanitizer where it is available."""
import sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import pointnet_numpy as onp
from tests import synth
from pointcloudprocessing_b200 import ops, _ffi
from pointcloudprocessing_b200.pointnet import PointNet
import os
os.environ["PCP_FPS_FORCE_PRUNE"] = "1"
# pruned FPS, small: cs = 1 and cs = 2 shapes, ties included
for b, n, s in ((3, 13000, 40), (2, 40000, 36), (2, 120000, 33)):
    x = synth.lidar_scans(9, b, n); x[:, : n // 3] = x[:, n // 3: 2 * (n // 3)]
    xd = torch.from_numpy(x).cuda()
    a = ops.farthest_point_sample(xd, s)
    r = torch.empty_like(a)
    _ffi.check(_ffi.load().pcp_fps_f32(_ffi.ptr(xd), b, n, s, 0, _ffi.ptr(r), _ffi.stream_ptr()))
    torch.cuda.synchronize()
    print("fps", b, n, bool(torch.equal(a, r)), ops.fps_workspace_bytes(b, n, s))
# fused tails + chains, small batch; staged host path
m = PointNet(23, 12, 0.3, 42, precision="bf16")
m.set_weights_dict(onp.make_params(23, 12, seed=7))
xh = synth.gaussian_clouds(3, 5, 256)
c1, _, R1 = m(torch.from_numpy(xh).cuda(), heads=("cls",))
c2, _, R2 = m(xh, heads=("cls",))
torch.cuda.synchronize()
print("model", bool(torch.equal(c1.cpu(), torch.as_tensor(c2))))
y = ops.normalize(torch.from_numpy(synth.gaussian_clouds(4, 3, 20000)).cuda())
torch.cuda.synchronize(); print("norm ok")
