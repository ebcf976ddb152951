"""Worker of tests/test_gpu_multigpu.py (launched with torchrun, one rank per GPU): data-parallel training step.
Checks on every rank that the all-reduced flat gradient bucket equals the sum of the per-shard gradients computed
locally without communication, that both ranks end with bit-identical weights after Adam, and that sharded inference
gathers to the single-GPU result."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from oracle import pointnet_numpy as onp  # noqa: E402
from pointcloudprocessing_b200 import parallel  # noqa: E402
from pointcloudprocessing_b200.pointnet import PointNet  # noqa: E402
from pointcloudprocessing_b200.pointnet import keras_compat as K  # noqa: E402
from tests import synth  # noqa: E402


def make_model(C, P):
    m = PointNet(C, P, 0.3, 42, precision="fp32")
    m.set_weights_dict(onp.make_params(C, P, seed=7))
    m.freeze_segmentation_head()
    m.compile(optimizer=K.Adam(learning_rate=K.ExponentialDecay(1e-3, 7000, 0.7)),
              loss_weights={"classification_output": 1.0, "segmentation_output": 0.0, "se3": 0.5})
    return m


def main():
    rank, world = parallel.init_from_env("nccl")
    C, P, gb, n = 7, 5, 8, 128
    rng = np.random.default_rng(5)
    x = synth.gaussian_clouds(5, gb, n)
    y = {"classification_output": rng.integers(0, C, gb).astype(np.int32),
         "segmentation_output": rng.integers(0, P, (gb, n)).astype(np.int32),
         "se3": np.tile(np.eye(3, dtype=np.float32), (gb, 1, 1))}
    xs, ys, off = parallel.shard_batch(x, y)
    m = make_model(C, P)
    m.train_on_batch(xs, ys, global_batch=gb, sample_offset=off, apply=False)          # all-reduced gradients
    reduced = m._grad_tensor().clone()
    ref = torch.zeros_like(reduced)
    m2 = make_model(C, P)
    for r in range(world):
        xr, yr, o = parallel.shard_batch(x, y, rank=r, world_size=world)
        m2.train_on_batch(xr, yr, global_batch=gb, sample_offset=o, apply=False, allreduce=False)
        ref += m2._grad_tensor()
    err = float((reduced - ref).abs().max() / ref.abs().max())
    assert err < 1e-5, err
    # one full step: identical weights on every rank
    m.train_on_batch(xs, ys, global_batch=gb, sample_offset=off)
    w = np.concatenate([v.ravel() for v in m.get_weights_dict().values() if v.size])
    digest = torch.tensor([float(np.abs(w).sum()), float((w * np.arange(w.size) % 7).sum())], dtype=torch.float64, device="cuda")
    lo, hi = digest.clone(), digest.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    # BN moving statistics are per shard (documented), everything Adam touches must agree: compare trainable tensors only
    wt = np.concatenate([v.ravel() for k, v in m.get_weights_dict().items() if "moving_" not in k])
    d2 = torch.tensor([float(np.abs(wt).sum())], dtype=torch.float64, device="cuda")
    a, b = d2.clone(), d2.clone()
    dist.all_reduce(a, op=dist.ReduceOp.MIN)
    dist.all_reduce(b, op=dist.ReduceOp.MAX)
    assert float(a) == float(b), (float(a), float(b))
    # sharded inference == whole-batch inference
    mi = PointNet(C, P, 0.3, 42, precision="bf16")
    mi.set_weights_dict(onp.make_params(C, P, seed=7))
    xi = synth.gaussian_clouds(6, 13, 256)
    xl, _, _ = parallel.shard_batch(xi)
    cls_l, _, _ = mi(torch.from_numpy(xl).cuda(), heads=("cls",))
    cls_all = parallel.gather_outputs(cls_l, len(xi))
    cls_ref, _, _ = mi(torch.from_numpy(xi).cuda(), heads=("cls",))
    assert torch.equal(cls_all, cls_ref)
    dist.barrier()
    if rank == 0:
        print(f"multigpu ok world={world} grad_err={err:.2e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
