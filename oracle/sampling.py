"""ORACLE — TEST INFRASTRUCTURE ONLY.  Definitions of farthest-point sampling and voxel-grid
downsampling, the two `north_star` sampling operators that have NO implementation in the
reference (the reference's only resize is truncate / random-duplicate,
PCA/pointcloud/PointCloudSet.py:443-470, restated in oracle/pointnet_numpy.py).

PARITY UNPINNED: there is nothing in /root/reference to pin these against.  This file *is* the
contract (SURVEY.md §8c "Oracle deliverables"); `oracle/sampling_c.c` is the same contract in
plain C (used for the timed CPU baseline) and tests check the two against each other and
against hand-derived known answers (cube corners, collinear ties, identical points, lattice).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32
VOXEL_KEY_BIAS = 1 << 20        # per-axis voxel index must satisfy |k| < 2^20


def fps(points: np.ndarray, s: int, start: int = 0) -> np.ndarray:
    """Farthest-point sampling of one cloud.  points (N,3) float32 -> (s,) int32 indices.

    d_0 = +inf; for each selected index i:  d = min(d, ((dx*dx + dy*dy) + dz*dz)) evaluated in fp32
    with separate roundings (no FMA contraction); next = argmax(d) with the LOWEST index winning
    ties (np.argmax).  The first index is `start`.  Once every distance is 0 the arg-max returns
    index 0 repeatedly (duplicates allowed).
    """
    p = np.ascontiguousarray(points, F32)
    n = p.shape[0]
    d = np.full(n, np.inf, F32)
    out = np.empty(s, np.int32)
    cur = int(start)
    x, y, z = p[:, 0].copy(), p[:, 1].copy(), p[:, 2].copy()
    for i in range(s):
        out[i] = cur
        dx = (x - x[cur]).astype(F32)
        dy = (y - y[cur]).astype(F32)
        dz = (z - z[cur]).astype(F32)
        dd = ((dx * dx).astype(F32) + (dy * dy).astype(F32)).astype(F32)
        dd = (dd + (dz * dz).astype(F32)).astype(F32)
        np.minimum(d, dd, out=d)
        cur = int(np.argmax(d))
    return out


def fps_batch(points: np.ndarray, s: int, start: int = 0) -> np.ndarray:
    return np.stack([fps(c, s, start) for c in points])


def voxel_keys(points: np.ndarray, origin: np.ndarray, voxel: float) -> np.ndarray:
    """k = floor((p - origin) / voxel) per axis: fp32 subtract, fp32 divide, floor -> int32."""
    p = np.asarray(points, F32)
    q = ((p - np.asarray(origin, F32)).astype(F32) / F32(voxel)).astype(F32)
    return np.floor(q).astype(np.int32)


def voxel_downsample(points: np.ndarray, voxel: float, origin=None):
    """Voxel-grid downsample of one cloud.  points (N,3) float32.

    origin=None -> the cloud's per-axis minimum.  Output voxels are ordered lexicographically by
    (kx, ky, kz).  Centroid = fp32( sum_in_float64(points in voxel) / count ) — double accumulation
    makes the result independent of summation order.  Returns (centroids (V,3) f32, counts (V,) i32,
    keys (V,3) i32).
    """
    p = np.ascontiguousarray(points, F32)
    if origin is None:
        origin = p.min(axis=0)
    k = voxel_keys(p, origin, voxel)
    if np.any(np.abs(k.astype(np.int64)) >= VOXEL_KEY_BIAS):
        raise ValueError("voxel index out of range (|k| >= 2^20)")
    kb = k.astype(np.int64) + VOXEL_KEY_BIAS
    lin = (kb[:, 0] << 42) | (kb[:, 1] << 21) | kb[:, 2]
    uniq, inv, counts = np.unique(lin, return_inverse=True, return_counts=True)
    sums = np.zeros((uniq.shape[0], 3), np.float64)
    np.add.at(sums, inv, p.astype(np.float64))
    cent = (sums / counts[:, None]).astype(F32)
    keys = np.stack([(uniq >> 42) & 0x1FFFFF, (uniq >> 21) & 0x1FFFFF, uniq & 0x1FFFFF], 1) - VOXEL_KEY_BIAS
    return cent, counts.astype(np.int32), keys.astype(np.int32)


def philox4x32_10(counter: np.ndarray, key: np.ndarray) -> np.ndarray:
    """Philox4x32-10 (Salmon et al. 2011), vectorised.  counter (...,4) uint32, key (...,2) uint32."""
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
    c = [counter[..., i].astype(np.uint32) for i in range(4)]
    k0 = np.broadcast_to(key[..., 0], c[0].shape).astype(np.uint32)
    k1 = np.broadcast_to(key[..., 1], c[0].shape).astype(np.uint32)
    for _ in range(10):
        p0 = c[0].astype(np.uint64) * M0
        p1 = c[2].astype(np.uint64) * M1
        hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), p0.astype(np.uint32)
        hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), p1.astype(np.uint32)
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + W0).astype(np.uint32)
        k1 = (k1 + W1).astype(np.uint32)
    return np.stack(c, -1)


def jitter_noise_bits(n_points: int, seed: int, offset: int = 0) -> np.ndarray:
    """The device jitter kernel's random bits: one Philox4x32-10 block per POINT, counter =
    (global point index + offset, 0, 0, 0) split lo/hi, key = (seed lo, seed hi).  Returns
    (n_points, 4) uint32.  Box-Muller on (u0,u1) gives the x,y normals; on (u2,u3) the z normal."""
    idx = np.arange(n_points, dtype=np.uint64) + np.uint64(offset)
    ctr = np.zeros((n_points, 4), np.uint32)
    ctr[:, 0] = (idx & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ctr[:, 1] = (idx >> np.uint64(32)).astype(np.uint32)
    key = np.array([seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF], np.uint32)
    return philox4x32_10(ctr, key)
