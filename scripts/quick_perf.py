"""Ad-hoc timings used while developing (not the bench contract): prints ms for a few kernels."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from pointcloudprocessing_b200 import ops
from tests import synth

def timeit(f, it=5, warm=2):
    for _ in range(warm): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it

x = torch.from_numpy(synth.lidar_scans(2002, 64, 100000)).cuda()
t = timeit(lambda: ops.farthest_point_sample(x, 1024), it=3, warm=1)
print(f"fps 64x100k->1024: {t:.2f} ms  {64*100000/t/1e3:.0f} Mpts/s  eff {64*1024*100000*20/t/1e6:.0f} GB/s")
t = timeit(lambda: ops.normalize(x))
print(f"normalize 64x100k: {t*1e3:.1f} us  {64*100000*24/t/1e6:.0f} GB/s")
sd = np.array([0.1,0.1,0.1], np.float32)
t = timeit(lambda: ops.jitter(x, sd, seed=1))
print(f"jitter 64x100k: {t*1e3:.1f} us  {64*100000*24/t/1e6:.0f} GB/s")
xs = torch.from_numpy(synth.lidar_scans(5005, 8, 131072)).cuda()
t = timeit(lambda: ops.voxel_downsample(xs, 0.35))
print(f"voxel 8x131072: {t*1e3:.1f} us")
if len(sys.argv) > 1:
    from oracle import pointnet_numpy as onp
    from pointcloudprocessing_b200.pointnet import PointNet
    for prec in sys.argv[1:]:
        m = PointNet(23, 12, 0.3, 42, precision=prec); m.set_weights_dict(onp.make_params(23, 12))
        for b in (32, 1024):
            xm = torch.from_numpy(synth.gaussian_clouds(1, b, 1024)).cuda()
            t = timeit(lambda: m(xm, heads=("cls",)))
            t2 = timeit(lambda: m(xm))
            print(f"{prec} b={b}: cls-only {t:.3f} ms ({b/t*1e3:.0f} clouds/s)  all {t2:.3f} ms ({b/t2*1e3:.0f} clouds/s)")
