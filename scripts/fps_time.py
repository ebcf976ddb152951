"""FPS wall time at BASELINE config-2 size (64 x 100 000 -> 1024): CUDA events, warm, 10 iterations."""
import sys; sys.path.insert(0, ".")
import torch
from pointcloudprocessing_b200 import ops
from tests import synth
x = torch.from_numpy(synth.lidar_scans(2002, 64, 100000)).cuda()
for _ in range(3): ops.farthest_point_sample(x, 1024)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): ops.farthest_point_sample(x, 1024)
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 10
print(f"fps 64x100000->1024: {t:.3f} ms  {64*100000/t/1e3:.0f} Mpts/s")
