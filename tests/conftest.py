import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib():
    """The C-ABI library; built in-tree if absent.  Never falls back to anything else."""
    from pointcloudprocessing_b200 import _ffi, build
    if not _ffi.lib_path().exists():
        build.build()
    return _ffi.load()


@pytest.fixture(scope="session")
def oracle_c():
    """oracle/_build/liboracle_sampling.so (plain-C restatement), built by oracle/Makefile."""
    import ctypes
    import subprocess
    so = ROOT / "oracle" / "_build" / "liboracle_sampling.so"
    if not so.exists():
        subprocess.check_call(["make", "-C", str(ROOT / "oracle")])
    return ctypes.CDLL(str(so))
