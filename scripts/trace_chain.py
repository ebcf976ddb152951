"""Development aid: run one bf16 forward with PCP_CHAIN_TRACE set and print the role timelines of cluster 0."""
import os, sys, glob
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ["PCP_CHAIN_TRACE"] = "/tmp/chain_trace"
import numpy as np, torch
from oracle import pointnet_numpy as onp
from pointcloudprocessing_b200.pointnet import PointNet
from tests import synth
m = PointNet(23, 12, 0.3, 42, precision="bf16"); m.set_weights_dict(onp.make_params(23, 12))
x = torch.from_numpy(synth.gaussian_clouds(1, 1024, 1024)).cuda()
for f in glob.glob("/tmp/chain_trace*"): os.remove(f)
m(x, heads=("cls",)); torch.cuda.synchronize()
for f in sorted(glob.glob("/tmp/chain_trace*"))[-3:]:
    a = np.fromfile(f, dtype=np.uint64).reshape(8, 2048, 2)
    t0 = min(a[s, 0, 1] for s in range(8) if a[s, 0, 0] != 0)
    print("====", f)
    for slot, name in [(0, "front0"), (2, "mma"), (1, "epi0"), (4, "front1"), (5, "epi1")]:
        ev = a[slot]
        n = int((ev[:, 0] != 0).sum())
        print(name, n, "events")
        # steady state: events of steps ~20..23
        rows = [(int(c), int(t - t0)) for c, t in ev[:n]]
        if name.startswith("front"):
            starts = [i for i, (c, t) in enumerate(rows) if c == 0x100]
            for si in starts[20:23]:
                seg = rows[si:si + 12]
                print("  ", " ".join(f"{c:x}@{t - seg[0][1]}" for c, t in seg), " step_start", seg[0][1])
        elif name == "mma":
            idx = [i for i, (c, t) in enumerate(rows) if c == 0x300]
            for si in idx[20:22]:
                seg = rows[si:si + 22]
                print("  ", " ".join(f"{c:x}@{t - seg[0][1]}" for c, t in seg), " t", seg[0][1])
        else:
            idx = [i for i, (c, t) in enumerate(rows) if c == 0x200]
            for si in idx[20:22]:
                seg = rows[si:si + 16]
                print("  ", " ".join(f"{c:x}@{t - seg[0][1]}" for c, t in seg), " t", seg[0][1])
