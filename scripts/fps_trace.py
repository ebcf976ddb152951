"""Per-warp timeline of the pruned FPS kernel (block 0, 16 iterations): PCP_FPS_PROF=1 python scripts/fps_trace.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloudprocessing_b200 import _ffi, ops  # noqa: E402
from tests import synth  # noqa: E402

assert os.environ.get("PCP_FPS_PROF"), "set PCP_FPS_PROF=1"
x = torch.from_numpy(synth.lidar_scans(2002, 64, 100_000)).cuda()
ops.farthest_point_sample(x, 1024)
buf = (C.c_longlong * (16 * 16 * 6))()
_ffi.check(_ffi.load().pcp_fps_trace(buf, 16 * 16 * 6))
t = np.array(buf, dtype=np.int64).reshape(16, 16, 6)
base = t[0, :, 0].min()
for it in range(16):
    r = t[it]
    top = r[:, 0]
    print(f"it {it}: loop top spread {top.min() - base:6d}..{top.max() - base:6d}  | per warp: dist (1-0), push (2-1), wake (5-2) ; reducer: xbar wake {r[0, 3] - base}, published {r[0, 4] - base}")
    print("    dist ", " ".join(f"{v:5d}" for v in (r[:, 1] - r[:, 0])))
    print("    push ", " ".join(f"{v:5d}" for v in (r[:, 2] - r[:, 1])))
    print("    pushed@", " ".join(f"{v:5d}" for v in (r[:, 2] - base)))
    print("    woke@", " ".join(f"{v:5d}" for v in (r[:, 5] - base)))
