from .PointCloudSet import PointCloudSet, get_dir_contents, load_from_file  # noqa: F401
