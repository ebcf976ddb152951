from .PointNet import PointNet, TNet, ConvLayer, DenseLayer, PointCloudNormalization  # noqa: F401
