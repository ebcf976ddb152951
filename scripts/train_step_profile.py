"""One BASELINE config-3 training step (classification_pretrain profile, 256 x 1024) for ncu launch lists / timing."""
import argparse, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
from oracle import pointnet_numpy as onp          # parameter init only (make_params); not on the timed path
from tests import synth
from pointcloudprocessing_b200 import ops
from pointcloudprocessing_b200.pointnet import PointNet, keras_compat as Kc

ap = argparse.ArgumentParser()
ap.add_argument("--precision", default="bf16")
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--points", type=int, default=1024)
ap.add_argument("--seg", action="store_true", help="segmentation profile (everything thawed, both losses)")
a = ap.parse_args()
C, P = 40, 50
dev = torch.device("cuda", 0)
tm = PointNet(C, P, 0.3, 42, precision=a.precision)
tm.set_weights_dict(onp.make_params(C, P, seed=7, trained_like=False))
tm.thaw_shared_network(); tm.thaw_input_transform(); tm.thaw_classification_head()
(tm.thaw_segmentation_head if a.seg else tm.freeze_segmentation_head)()
tm.compile(optimizer=Kc.Adam(learning_rate=Kc.ExponentialDecay(1e-4, 7000, 0.7)),
           loss_weights={"classification_output": 1.0, "segmentation_output": 1.0 if a.seg else 0.0, "se3": 0.0})
rng = np.random.default_rng(3003)
xt = torch.from_numpy(synth.gaussian_clouds(3003, a.batch, a.points)).to(dev)
yt = {"classification_output": torch.from_numpy(rng.integers(0, C, a.batch).astype(np.int32)).to(dev),
      "segmentation_output": torch.from_numpy(rng.integers(0, P, (a.batch, a.points)).astype(np.int32)).to(dev),
      "se3": torch.eye(3, device=dev).repeat(a.batch, 1, 1)}
for i in range(a.steps):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    met = tm.train_on_batch(ops.jitter(xt, (0.1, 0.1, 0.1), seed=i), yt)
    torch.cuda.synchronize()
    print("step", i, "%.2f ms" % ((time.perf_counter() - t0) * 1e3), "loss", met["loss"], flush=True)
