"""pointcloudprocessing_b200 — B200 (sm_100a) implementation of the PointNet point-cloud hot path of
MAPieschl/PointCloudProcessing behind the reference's own Python API (pointnet/, pointcloud/)."""
__version__ = "0.1.0"
