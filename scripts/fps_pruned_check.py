"""Pruned FPS (pcp_fps_ws_f32) against the unpruned kernel (pcp_fps_f32) and the C oracle: bit-exactness on awkward clouds, timing,
and the fraction of (warp, group) updates actually executed (PCP_FPS_PROF=1).  Run on the GPU box: python scripts/fps_pruned_check.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloudprocessing_b200 import _ffi, ops  # noqa: E402
from tests import synth  # noqa: E402


def plain(x, s, start=0):
    b, n, _ = x.shape
    idx = torch.empty((b, s), device=x.device, dtype=torch.int32)
    _ffi.check(_ffi.load().pcp_fps_f32(_ffi.ptr(x), b, n, s, start, _ffi.ptr(idx), _ffi.stream_ptr()))
    return idx


def timed(f, reps=5):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    ts = []
    for _ in range(reps):
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def cases():
    rng = np.random.default_rng(5)
    yield "lidar 64x100000", synth.lidar_scans(2002, 64, 100_000), 1024, 0
    yield "lidar 64x70001 start 77", synth.lidar_scans(11, 64, 70_001), 512, 77
    x = synth.lidar_scans(12, 40, 100_000)
    x[:, 50_000:] = x[:, :50_000]                        # every point twice: ties on every arg-max
    yield "duplicated halves 40x100000", x, 256, 3
    x = np.round(rng.uniform(-4, 4, (40, 90_000, 3))).astype(np.float32)    # integer lattice: masses of exact ties
    yield "lattice 40x90000", x, 300, 0
    x = np.zeros((40, 80_000, 3), np.float32) + 1.5
    yield "all identical 40x80000", x, 64, 5
    x = rng.normal(0, 1, (48, 100_000, 3)).astype(np.float32) * np.float32(1e-3) + np.float32(1000.0)   # tiny cloud far from the origin
    yield "offset 48x100000", x, 400, 0
    x = rng.uniform(-50, 50, (64, 100_000, 3)).astype(np.float32)
    yield "uniform cube 64x100000", x, 1024, 0


def main():
    prof = os.environ.get("PCP_FPS_PROF") is not None
    lib = _ffi.load()
    for name, x, s, start in cases():
        xd = torch.from_numpy(x).cuda()
        b, n, _ = x.shape
        ws = ops.fps_workspace_bytes(b, n, s)
        a = ops.farthest_point_sample(xd, s, start)
        r = plain(xd, s, start)
        same = bool(torch.equal(a, r))
        t_new = timed(lambda: ops.farthest_point_sample(xd, s, start))
        t_old = timed(lambda: plain(xd, s, start))
        extra = ""
        if prof and ws:
            ops.farthest_point_sample(xd, s, start)
            out = (C.c_ulonglong * 8)()
            lib.pcp_fps_phase_cycles(out)
            ph = list(out)
            extra = (f" cycles/iter dist {ph[0] / (s - 1):.0f} sync {ph[2] / (s - 1):.0f} push {ph[3] / (s - 1):.0f} wait {ph[4] / (s - 1):.0f}"
                     f" read {ph[5] / (s - 1):.0f}; groups updated {ph[6] / (16 * 7 * (s - 1)):.3f}; tie iterations {ph[7]} of {s - 1}")
        print(f"{name}: ws {ws / 1e6:.1f} MB identical {same}  pruned {t_new:.3f} ms  plain {t_old:.3f} ms  x{t_old / t_new:.2f}{extra}", flush=True)
        if not same:
            bad = (a != r).any(1).nonzero().flatten()[:4].tolist()
            c = bad[0]
            j = int((a[c] != r[c]).nonzero()[0])
            print(f"   MISMATCH clouds {bad}: cloud {c} first at sample {j}: {int(a[c, j])} vs {int(r[c, j])}")


if __name__ == "__main__":
    main()
