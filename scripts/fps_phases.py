"""Where an FPS iteration goes (64 x 100 000 -> 1024, cluster of 8, 25 points per thread): PCP_FPS_PROF=1 makes block 0 /
thread 0 accumulate clock64 deltas per phase; printed as cycles per iteration."""
import ctypes as C, os, sys
os.environ["PCP_FPS_PROF"] = "1"
sys.path.insert(0, ".")
import torch
from pointcloudprocessing_b200 import ops, _ffi
from tests import synth
x = torch.from_numpy(synth.lidar_scans(2002, 64, 100000)).cuda()
ops.farthest_point_sample(x, 1024); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ops.farthest_point_sample(x, 1024); e1.record(); torch.cuda.synchronize()
if os.environ.get("PCP_FPS_TIME_ONLY"):
    print("total %.2f ms (profiling stamps on)" % e0.elapsed_time(e1)); sys.exit(0)
out = (C.c_ulonglong * 8)()
lib = _ffi.load()
_ffi.check(lib.pcp_fps_phase_cycles(out))
names = ["distance loop", "candidate search + record", "__syncthreads", "warp-0 stage + pushes", "cluster barrier", "slot read-back", "-", "-"]
print("total %.2f ms" % e0.elapsed_time(e1))
for nm, v in zip(names, out):
    print(f"{nm:28s} {v / 1023:8.0f} cycles / iteration")
print(f"{'sum':28s} {sum(out) / 1023:8.0f}")
