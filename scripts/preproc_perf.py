"""Timings of the streaming / sampling kernels at BASELINE config-2 size (64 x 100 000 points): CUDA events, warm, 20 iterations.
Used for ncu launch lists too (`ncu --metrics gpu__time_duration.sum ... python scripts/preproc_perf.py`)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch
from pointcloudprocessing_b200 import ops
from tests import synth

HBM = 6454.3


def timeit(f, it=20, warm=3):
    for _ in range(warm): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / it


b, n = 64, 100_000
x = torch.from_numpy(synth.lidar_scans(2002, b, n)).cuda()
t = timeit(lambda: ops.normalize(x))
print(f"normalize {b}x{n}: {t*1e3:.1f} us  {b*n*24/t/1e6:.0f} GB/s  {b*n*24/t/1e6/HBM:.3f} of HBM")
sd = np.array([0.1, 0.1, 0.1], np.float32)
t = timeit(lambda: ops.jitter(x, sd, seed=1))
print(f"jitter {b}x{n}: {t*1e3:.1f} us  {b*n*24/t/1e6:.0f} GB/s  {b*n*24/t/1e6/HBM:.3f} of HBM")
cent, cnt, nv = ops.voxel_downsample(x, 0.7)
V = int(nv.sum())
t = timeit(lambda: ops.voxel_downsample(x, 0.7))
by = 12 * b * n + 16 * V
print(f"voxel {b}x{n} @0.7m (V/cloud {V/b:.0f}): {t*1e3:.1f} us  {b*n/t/1e3:.0f} Mpts/s  {by/t/1e6:.0f} GB/s  {by/t/1e6/HBM:.3f} of HBM (12N+16V)")
