"""End-to-end scan pipeline (BASELINE config 5): raw LiDAR-sized scans -> voxel-grid downsample -> resize to the
network width (PointCloudSet._adjust_to_input_width semantics, pointcloud/PointCloudSet.py:443-470: keep the first
`width` points / pad with random duplicates) -> PointNet (normalisation is the model's first layer,
pointnet/PointNet.py:202).  Every stage is a libpcp_b200.so kernel; tensors stay on the device between stages."""
from __future__ import annotations

import torch

from . import ops


def voxel_resize(scans: torch.Tensor, voxel: float, width: int, seed: int = 0) -> torch.Tensor:
    """(b, N_raw, 3) CUDA scans -> (b, width, 3) voxel centroids, truncated / duplicate-padded per cloud."""
    cen, _cnt, nvox = ops.voxel_downsample(scans, voxel, capacity=width)          # only the first `width` voxels are ever read
    return ops.resize_to_width(cen, nvox, width, seed=seed)


def classify_scans(model, scans: torch.Tensor, voxel: float, width: int, seed: int = 0, heads=("cls",)):
    """Returns the model outputs [cls (b,C), seg or None, R (b,3,3)] for raw scans."""
    return model(voxel_resize(scans, voxel, width, seed), heads=heads)
