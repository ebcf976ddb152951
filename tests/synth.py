"""Synthetic inputs of SURVEY.md §8(d) (seeded, float32, C-contiguous)."""
import numpy as np


def gaussian_clouds(seed, b, n):
    """C1/C3/C4 generator: per cloud centre U(-50,50)^3 m, per-axis scale U(5,30) m, points centre + scale*N(0,1)."""
    rng = np.random.default_rng(seed)
    centre = rng.uniform(-50, 50, (b, 1, 3))
    scale = rng.uniform(5, 30, (b, 1, 3))
    return np.ascontiguousarray((centre + scale * rng.standard_normal((b, n, 3))).astype(np.float32))


def lidar_scans(seed, b, n, dup_frac=0.01):
    """C2/C5 generator: mixture of 6 anisotropic Gaussians (aircraft-like) at 50-300 m range along +x,
    quantised to 1 mm, plus `dup_frac` exact duplicates to exercise distance ties."""
    rng = np.random.default_rng(seed)
    parts = np.array([[30, 3, 3], [3, 15, 0.5], [3, 15, 0.5], [2, 6, 0.4], [2, 0.4, 4], [3, 1.5, 1.5]], np.float64) / 2.0
    offs = np.array([[0, 0, 0], [0, 9, 0], [0, -9, 0], [-13, 0, 0.5], [-13, 0, 2.5], [3, 6, -1.5]], np.float64)
    w = np.array([0.4, 0.15, 0.15, 0.1, 0.1, 0.1])
    out = np.empty((b, n, 3), np.float32)
    for c in range(b):
        comp = rng.choice(6, size=n, p=w)
        pts = offs[comp] + parts[comp] * rng.standard_normal((n, 3))
        pts[:, 0] += rng.uniform(50, 300)
        pts = np.round(pts, 3)
        nd = int(n * dup_frac)
        if nd:
            src = rng.integers(0, n, nd)
            dst = rng.integers(0, n, nd)
            pts[dst] = pts[src]
        out[c] = pts.astype(np.float32)
    return np.ascontiguousarray(out)
