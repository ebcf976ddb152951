"""Development aid: time ops.VoxelGrid (pcp_voxel_compact_f32) against pcp_voxel_f32 at config-2/5 shapes."""
import sys
from pathlib import Path
import numpy as np
import torch
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from pointcloudprocessing_b200 import ops  # noqa: E402
from tests import synth  # noqa: E402


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


for b, n, voxel, cap in ((64, 100_000, 0.7, 16384), (8, 131072, 0.7, 16384), (64, 100_000, 0.35, 65536)):
    x = torch.from_numpy(synth.lidar_scans(2002, b, n)).cuda()
    vg = ops.VoxelGrid(b, n, cap)
    ms = timed(lambda: vg(x, voxel))
    nv = vg.nvox.cpu().numpy()
    alg = 12 * b * n + 16 * int(np.clip(nv, 0, None).sum())
    ms_old = timed(lambda: ops._voxel_general(x, voxel, None, n), iters=5)
    print(f"b={b} n={n} voxel={voxel}: compact {ms*1e3:.1f} us ({alg/ms/1e6:.0f} GB/s of 12N+16V, V/cloud {nv.mean():.0f}) | pcp_voxel_f32 {ms_old*1e3:.1f} us (L2 flushed before every call)")
