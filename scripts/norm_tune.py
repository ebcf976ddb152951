"""normalize at 64 x 100 000 for one slice size (PCP_NORM_SLICE_KB, read once per process).
PCP_NORM_DBG bits (timing experiments only, results are wrong): 1 = no waits on the cloud counters, 2 = no sum / max loops,
4 = no write loop.  profiles/r01_norm_tune.txt holds the table."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from pointcloudprocessing_b200 import ops
from tests import synth
x = torch.from_numpy(synth.lidar_scans(2002, 64, 100000)).cuda()
for _ in range(3): ops.normalize(x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): ops.normalize(x)
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 50
print(os.environ.get("PCP_NORM_SLICE_KB", "default"), f"KB: {t*1e3:.1f} us  {64*100000*24/t/1e6:.0f} GB/s")
