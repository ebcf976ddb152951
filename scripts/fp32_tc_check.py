"""Development aid / profiles: the three inference precisions against the oracle (max |err| / max |ref| on the class logits, R and
the segmentation logits) and their throughput at 1024 x 1024 points, cls + R."""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from oracle import pointnet_numpy as onp  # noqa: E402
from oracle import pointnet_torch as opt  # noqa: E402
from pointcloudprocessing_b200.pointnet import PointNet  # noqa: E402
from tests import synth  # noqa: E402


def rel(a, r):
    return float(np.abs(a.astype(np.float64) - r).max() / np.abs(r).max())


p = onp.make_params(23, 12, seed=7)
x = synth.gaussian_clouds(1001, 32, 1024)
with torch.no_grad():
    ref = {k: v.numpy() for k, v in opt.forward(torch.from_numpy(x).double(), {k: v.double() for k, v in opt.to_torch(p).items()}).items()} \
        if False else {k: v.numpy() for k, v in opt.forward(torch.from_numpy(x), opt.to_torch(p)).items()}
xb = torch.from_numpy(np.concatenate([synth.gaussian_clouds(5, 512, 1024)] * 2)).cuda()
for prec in ("fp32", "fp32_tc", "bf16"):
    m = PointNet(23, 12, 0.3, 42, precision=prec)
    m.set_weights_dict(p)
    out = m.forward_with_logits(torch.from_numpy(x).cuda())
    errs = {k: rel(out[k].cpu().numpy(), ref[k]) for k in ("R", "cls_logits", "seg_logits")}
    for _ in range(2):
        m(xb, heads=("cls",))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        m(xb, heads=("cls",))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(f"{prec:8s} err vs oracle: " + ", ".join(f"{k} {v:.2e}" for k, v in errs.items()) + f" | 1024 clouds x 1024 pts (cls + R): {ms:.2f} ms = {1024 / ms * 1e3:,.0f} clouds/s")
