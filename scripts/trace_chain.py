"""Development aid: run one bf16 forward with PCP_CHAIN_TRACE set and print, for the third chain launch,
the per-granule pipeline timeline of cluster 0 (ns): MMA issue -> acc_full seen by both CTAs' epilogues
-> acc_empty arrivals, plus the front-chain events."""
import glob
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
os.environ["PCP_CHAIN_TRACE"] = "/tmp/chain_trace"
import numpy as np
import torch
from oracle import pointnet_numpy as onp
from pointcloudprocessing_b200.pointnet import PointNet
from tests import synth

m = PointNet(23, 12, 0.3, 42, precision="bf16")
m.set_weights_dict(onp.make_params(23, 12))
x = torch.from_numpy(synth.gaussian_clouds(1, 1024, 1024)).cuda()
for f in glob.glob("/tmp/chain_trace*"):
    os.remove(f)
m(x, heads=("cls",))
torch.cuda.synchronize()
which = int(sys.argv[1]) if len(sys.argv) > 1 else -1          # launch of the forward: 0 / 1 / 2 = chain A / B / C (default C)
f = sorted(glob.glob("/tmp/chain_trace*"))[which]
a = np.fromfile(f, dtype=np.uint64).reshape(8, 2048, 2)


def events(slot):
    ev = a[slot]
    n = int((ev[:, 0] != 0).sum())
    return [(int(c), int(t)) for c, t in ev[:n]]


mma, fin, e0, e1, f0, f1 = events(2), events(3), events(1), events(5), events(0), events(4)
t0 = mma[0][1]
# granules in issue order: final events 0x400 + hf*16 + j*2 + kb ; a granule completes at kb == 1
gr = [[c, t, t] for c, t in fin if c >= 0x400]
seen0 = [t for c, t in e0 if 0x200 <= c < 0x210]
arr0 = [t for c, t in e0 if 0x210 <= c < 0x220]
seen1 = [t for c, t in e1 if 0x200 <= c < 0x210]
arr1 = [t for c, t in e1 if 0x210 <= c < 0x220]
print("granule  issue_kb0  issue_kb1(commit)  seen0  seen1  arrived0  arrived1   (ns rel. to first MMA event); next-use issue of same buffer")
for g in range(160, min(184, len(gr) - 2)):
    nxt = gr[g + 2][1] - t0 if g + 2 < len(gr) else -1
    print(f"g={g:4d} code={gr[g][0]:x}  {gr[g][1]-t0:8d} {gr[g][2]-t0:8d}   {seen0[g]-t0:8d} {seen1[g]-t0:8d}   {arr0[g]-t0:8d} {arr1[g]-t0:8d}    next {nxt:8d}")
print("mid issues (code, t):", [(hex(c), t - t0) for c, t in mma if c < 0x400][60:72])
W = 200000
def win(ev, lo=W, hi=W + 9000):
    return [(hex(c), t - t0) for c, t in ev if lo <= t - t0 < hi]
print("WINDOW front0:", win(f0))
print("WINDOW front1:", win(f1))
print("WINDOW final :", win(fin))
print("WINDOW mma   :", win(mma))
print("front0:", [(hex(c), t - t0) for c, t in f0][220:245])
print("front1:", [(hex(c), t - t0) for c, t in f1][220:245])
