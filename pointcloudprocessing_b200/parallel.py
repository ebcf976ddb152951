"""Multi-GPU plumbing: one process per GPU, contiguous batch shards, no data-path collective for
inference / preprocessing / sampling; training adds ONE all-reduce over the flat gradient bucket.

The reference is single-process (pointnet_train.py:385-388 only enables memory growth on GPU 0); this module is new
design, not parity work.  torch.distributed (NCCL over NVLink on the GPU box, gloo in the CPU tests) is the transport.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def world():
    """(rank, world_size) — (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def barrier():
    """dist.barrier() when a process group is up; nothing otherwise (single-process runs)."""
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def bind_to_gpu_numa(device_index: int | None = None) -> dict:
    """Pin this process to the CPU cores NVML reports as local to its GPU (the GPU's NUMA node), BEFORE any pinned host buffer is
    allocated: pinned pages are first-touch, so the end-to-end path's H2D copies then start from the memory controller next to the
    GPU's PCIe root instead of wherever the launcher happened to start the rank (eight ranks pulling 50 MB per
    step through one node).  Returns what was done — {"bound": bool, "cpus": n, "reason": ...} — and never raises: on a box whose
    GPUs all report the same CPU set (one NUMA node) there is nothing to gain and the mask is left alone."""
    info = {"bound": False, "cpus": 0, "reason": ""}
    try:
        import pynvml
        if device_index is None:
            device_index = torch.cuda.current_device() if torch.cuda.is_available() else 0
        pynvml.nvmlInit()
        n_dev = pynvml.nvmlDeviceGetCount()
        words = (os.cpu_count() + 63) // 64

        def cpus_of(i):
            # honour CUDA_VISIBLE_DEVICES: torch's ordinal i is the i-th visible device
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = i
            if vis:
                ids = [v.strip() for v in vis.split(",") if v.strip()]
                if i < len(ids) and ids[i].isdigit():
                    phys = int(ids[i])
            h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
            return {w * 64 + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        mine = cpus_of(device_index)
        allowed = os.sched_getaffinity(0)
        target = mine & allowed
        info["cpus"] = len(target)
        if not target:
            info["reason"] = "no overlap between the GPU's CPU set and this process's allowed CPUs"
        elif all(cpus_of(i) == mine for i in range(n_dev)) and n_dev > 1:
            info["reason"] = "every GPU reports the same CPU set (single NUMA node): nothing to bind"
        elif target == allowed:
            info["reason"] = "already confined to the GPU's CPU set"
        else:
            os.sched_setaffinity(0, target)
            info["bound"] = True
    except Exception as e:                      # NVML missing, permission denied, exotic topology: keep running unbound
        info["reason"] = f"{type(e).__name__}: {e}"
    return info


def init_from_env(backend: str | None = None):
    """Initialise the default process group from torchrun's RANK / WORLD_SIZE / MASTER_* environment."""
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    if ws <= 1 or dist.is_initialized():
        return world()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        lr = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(lr)
        bind_to_gpu_numa(lr)
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    else:
        dist.init_process_group(backend)
    return world()


def shard_bounds(total: int, rank: int, world_size: int):
    """Contiguous shard [lo, hi) of `total` units for `rank`: sizes differ by at most one, earlier ranks take the extra."""
    if not (0 <= rank < world_size):
        raise ValueError("rank out of range")
    base, extra = divmod(total, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(x, labels: dict | None = None, rank: int | None = None, world_size: int | None = None):
    """This rank's slice of a global batch (array-likes indexed on axis 0) and its global sample offset, which keys the
    Philox streams (jitter, dropout) so that results do not depend on the number of shards."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    lo, hi = shard_bounds(len(x), rank, world_size)
    ys = None if labels is None else {k: v[lo:hi] for k, v in labels.items()}
    return x[lo:hi], ys, lo


def allreduce_gradients(flat: torch.Tensor):
    """SUM all-reduce of the flat gradient bucket.  The loss kernel already divides by the GLOBAL batch, so the sum over
    ranks is the global-mean gradient and every rank applies the identical Adam step."""
    if world()[1] > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    return flat


def gather_outputs(local: torch.Tensor, total: int):
    """Concatenate per-shard outputs in rank order on every rank (inference results; shards may differ by one row)."""
    r, w = world()
    if w == 1:
        return local
    sizes = [hi - lo for lo, hi in (shard_bounds(total, i, w) for i in range(w))]
    rows = max(sizes)                                    # all_gather wants equal shapes: pad the short shards by one row
    padded = torch.zeros((rows,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    bufs = [torch.empty_like(padded) for _ in range(w)]
    dist.all_gather(bufs, padded)
    return torch.cat([b[:k] for b, k in zip(bufs, sizes)], 0)


def max_over_ranks(value: float, device=None) -> float:
    """Timing reduction of bench.py: the slowest rank defines the step."""
    r, w = world()
    if w == 1:
        return float(value)
    t = torch.tensor([float(value)], device=device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
