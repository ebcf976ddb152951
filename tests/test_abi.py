"""CPU: the C-ABI library loads and exports exactly what include/pcp.h declares; no CPU fallback."""
import ctypes
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]


def test_header_and_binding_agree(lib):
    from pointcloudprocessing_b200 import _ffi
    # pcp.h: the drop-in boundary; pcp_dev.h: test / development hooks (kept out of the boundary header)
    hdr = (ROOT / "include" / "pcp.h").read_text() + (ROOT / "include" / "pcp_dev.h").read_text()
    declared = set(re.findall(r"\b(pcp_[a-z0-9_]+)\s*\(", hdr))
    product = set(re.findall(r"\b(pcp_[a-z0-9_]+)\s*\(", (ROOT / "include" / "pcp.h").read_text()))
    assert not [n for n in product if "selftest" in n or "bench" in n or "phase_cycles" in n], "test hooks belong in pcp_dev.h"
    assert declared == set(_ffi.SIGNATURES), (declared ^ set(_ffi.SIGNATURES))
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in pcp.h but not exported"


def test_version_and_error_string(lib):
    assert lib.pcp_version() >= 100
    assert isinstance(lib.pcp_last_error(), bytes)


def test_product_never_imports_oracle():
    pkg = ROOT / "pointcloudprocessing_b200"
    for f in pkg.rglob("*.py"):
        txt = f.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, re.M), f"{f} imports the oracle"
        assert "from .. import oracle" not in txt and "import oracle" not in txt.replace("oracle/", "")
    for f in pkg.rglob("*.cu*"):
        assert "#include \"../../oracle" not in f.read_text()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu(lib):
    from pointcloudprocessing_b200 import _ffi
    desc = _ffi.ModelDesc(23, 12, 0, 0, 0, 0.3, _ffi.PRECISION_FP32)
    h = ctypes.c_void_p()
    st = lib.pcp_model_create(ctypes.byref(desc), ctypes.byref(h))
    assert st == -2, "model creation must fail with PCP_ERR_CUDA when there is no device"
    assert b"no CPU fallback" in lib.pcp_last_error()


def test_argument_validation_without_gpu(lib):
    assert lib.pcp_normalize_f32(None, 1, 10, None, None, None, None) == -1
    assert lib.pcp_fps_f32(None, 1, 10, 4, 0, None, None) == -1
    n = ctypes.c_size_t()
    assert lib.pcp_voxel_workspace_bytes(2, 1000, ctypes.byref(n)) == 0 and n.value > 2 * 1000 * 24
