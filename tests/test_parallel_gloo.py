"""CPU, world_size 2, gloo: the host-side multi-GPU logic of pointcloudprocessing_b200/parallel.py — shard arithmetic,
the gradient-bucket all-reduce contract (sum of per-shard gradients scaled by 1/global_batch == whole-batch gradient),
rank-ordered output gather and max-over-ranks timing.  The per-shard arithmetic itself is GPU work (tests -m gpu)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pointcloudprocessing_b200 import parallel


def test_shard_bounds_cover_without_overlap():
    for total in (0, 1, 7, 8, 32, 255, 256):
        for w in (1, 2, 3, 4, 8):
            spans = [parallel.shard_bounds(total, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        parallel.shard_bounds(8, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    assert parallel.init_from_env("gloo") == (rank, world)
    rng = np.random.default_rng(0)
    gb, d = 9, 33                                         # odd global batch: shards of 5 and 4
    x = rng.standard_normal((gb, d)).astype(np.float32)
    y = {"classification_output": np.arange(gb, dtype=np.int32)}
    w = torch.from_numpy(rng.standard_normal(d).astype(np.float32))
    xs, ys, off = parallel.shard_batch(x, y)
    lo, hi = parallel.shard_bounds(gb, rank, world)
    assert off == lo and len(xs) == hi - lo and np.array_equal(ys["classification_output"], np.arange(lo, hi))
    # toy loss  L = sum_i (x_i . w)^2 / global_batch : per-shard gradient already carries 1/global_batch (like pcp_model_loss_backward)
    xt = torch.from_numpy(xs)
    flat = (2.0 * (xt @ w)[:, None] * xt).sum(0) / gb
    parallel.allreduce_gradients(flat)
    whole = (2.0 * (torch.from_numpy(x) @ w)[:, None] * torch.from_numpy(x)).sum(0) / gb
    ok_grad = bool(torch.allclose(flat, whole, rtol=1e-5, atol=1e-6))
    out = parallel.gather_outputs(torch.from_numpy(xs) * 2.0, gb)
    ok_gather = bool(torch.equal(out, torch.from_numpy(x) * 2.0))
    ok_max = parallel.max_over_ranks(1.0 + rank) == float(world)
    ret[rank] = (ok_grad, ok_gather, ok_max)
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert dict(ret) == {0: (True, True, True), 1: (True, True, True)}


def test_single_process_defaults():
    assert parallel.world() == (0, 1)
    t = torch.ones(4)
    assert parallel.allreduce_gradients(t) is t and parallel.max_over_ranks(3.5) == 3.5
    x = np.arange(10)
    xs, ys, off = parallel.shard_batch(x, None, rank=1, world_size=4)
    assert list(xs) == [3, 4, 5] and ys is None and off == 3


def test_bind_to_gpu_numa_never_raises():
    """Host-side placement helper: without NVML / a GPU it reports why it did nothing and leaves the CPU mask alone."""
    import os
    from pointcloudprocessing_b200 import parallel
    before = os.sched_getaffinity(0)
    info = parallel.bind_to_gpu_numa(0)
    assert set(info) == {"bound", "cpus", "reason"}
    if not info["bound"]:
        assert os.sched_getaffinity(0) == before
    else:
        os.sched_setaffinity(0, before)
