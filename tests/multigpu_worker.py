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


def make_model(C, P, precision="fp32", sync_bn=False, smooth=False):
    m = PointNet(C, P, 0.3, 42, precision=precision)
    p = onp.make_params(C, P, seed=7)
    if smooth:                                  # every ReLU active: the gradient is a continuous function of round-off
        p = {k: ((v + np.float32(6.0)).astype(np.float32) if k.endswith("/bn/beta") else v) for k, v in p.items()}
    m.set_weights_dict(p)
    m.freeze_segmentation_head()
    m.compile(optimizer=K.Adam(learning_rate=K.ExponentialDecay(1e-3, 7000, 0.7)),
              loss_weights={"classification_output": 1.0, "segmentation_output": 0.0, "se3": 0.5}, sync_bn=sync_bn)
    return m


def check_sync_bn(x, y, gb, C, P, world, precision, tol):
    """sync_bn: the data-parallel step (world x gb/world clouds, BatchNorm sums all-reduced inside the step) equals the
    single-device step on the gb clouds — loss, all-reduced gradients and BN moving statistics."""
    xs, ys, off = parallel.shard_batch(x, y)
    md = make_model(C, P, precision, sync_bn=True, smooth=True)
    met_d = md.train_on_batch(xs, ys, global_batch=gb, sample_offset=off, apply=False)
    g_dp = md._grad_tensor().clone()
    ms = make_model(C, P, precision, sync_bn=False, smooth=True)
    met_s = ms.train_on_batch(x, y, global_batch=gb, sample_offset=0, apply=False, allreduce=False)
    g_1 = ms._grad_tensor().clone()
    err = float((g_dp - g_1).abs().max() / g_1.abs().max())
    cos = float((g_dp * g_1).sum() / (g_dp.norm() * g_1.norm()))
    if os.environ.get("PCP_MG_REPORT") and dist.get_rank() == 0:
        print(f"[report] {precision}: max-norm err {err:.3e} rel-L2 {float((g_dp - g_1).norm() / g_1.norm()):.3e} cosine {cos:.5f}", flush=True)
    loss = torch.tensor([met_d["classification_output_loss"] * len(xs) / gb], dtype=torch.float64, device="cuda")
    dist.all_reduce(loss)          # the per-rank loss is this shard's share of the global mean: summed over the ranks it is the single-device loss
    loss_1 = met_s["classification_output_loss"]
    if precision == "fp32":
        assert err < tol, (precision, err)
        assert abs(float(loss) - loss_1) <= 1e-4 * max(1.0, loss_1)
    else:
        # bf16: the pooled layers take their batch statistics from the Gram matrix of their input, and one device sums it in a
        # different order than two (per-rank sums, then the all-reduce) — a 1e-7 difference that a bf16-rounded network amplifies
        # (test_gpu_train.py docstring): on 2 x B200 the two losses of this batch are 8.414 and 8.509, the gradients' cosine 0.95.
        # With PCP_TRAIN_EXPLICIT_POOL=1 the comparison agrees to 3e-7, the pooled layer's protocol is pinned exactly on one device
        # (test_gpu_train_pool.py::test_pooled_layer_data_parallel_protocol) and end to end in fp32 above; here direction, scale and
        # the loss are bounded.  Each side is reproducible by itself: the Gram matrix is summed in a fixed order (tc_gram_det), so
        # the single-device step run twice gives the same loss bit for bit — checked below (before round 2's fix it varied by 1 %
        # from run to run and this test failed two times in five).
        ms2 = make_model(C, P, precision, sync_bn=False, smooth=True)
        met_s2 = ms2.train_on_batch(x, y, global_batch=gb, sample_offset=0, apply=False, allreduce=False)
        g_1b = ms2._grad_tensor().clone()
        self_cos = float((g_1b * g_1).sum() / (g_1b.norm() * g_1.norm()))
        self_loss = abs(met_s2["classification_output_loss"] - loss_1) / max(1.0, loss_1)
        if os.environ.get("PCP_MG_REPORT") and dist.get_rank() == 0:
            print(f"[report] {precision}: single-device step against itself: cosine {self_cos:.5f} loss rel diff {self_loss:.3e}; "
                  f"data-parallel loss {float(loss):.6f} single {loss_1:.6f}", flush=True)
        assert self_loss <= 1e-6 and self_cos > 0.999, ("the bf16 step is not reproducible run to run", self_loss, self_cos)
        assert cos > 0.8 and 0.5 < float(g_dp.norm() / g_1.norm()) < 2.0, (precision, err, cos, self_cos)
        assert abs(float(loss) - loss_1) / max(1.0, loss_1) <= 3e-2, (float(loss), loss_1, self_loss)
    wd, w1 = md.get_weights_dict(), ms.get_weights_dict()
    for k in wd:
        if "moving_" in k and precision == "fp32":
            ref = np.abs(w1[k]).max() + 1e-12
            assert np.abs(wd[k] - w1[k]).max() / ref < 1e-4, (precision, k)
    return err


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
    # data-parallel BatchNorm: equal shards required
    gb2 = 8 * world
    x2 = synth.gaussian_clouds(15, gb2, n)
    y2 = {"classification_output": rng.integers(0, C, gb2).astype(np.int32), "segmentation_output": rng.integers(0, P, (gb2, n)).astype(np.int32),
          "se3": np.tile(np.eye(3, dtype=np.float32), (gb2, 1, 1))}
    e32 = check_sync_bn(x2, y2, gb2, C, P, world, "fp32", 5e-4)
    e16 = check_sync_bn(x2, y2, gb2, 8, 8, world, "bf16", None)
    dist.barrier()
    if rank == 0:
        print(f"multigpu ok world={world} grad_err={err:.2e} sync_bn fp32 {e32:.2e} bf16 {e16:.2e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
