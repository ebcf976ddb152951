"""pcp_normalize_noalloc_f32 — the allocation-free variant of the sliced normalisation kernel (scratch in a static device
array; no counterpart in the reference).  Called through the C-ABI and checked
against the existing entry point pcp_normalize_f32 on two calls in a row (the second with fewer clouds, so stale scratch
contents from the first would show), plus the size limit.  Collected last (file name) so a failure here hides nothing."""
import ctypes as C

import numpy as np
import pytest
import torch

from tests import synth

pytestmark = pytest.mark.gpu


def _ulp_err(a, ref, scale):
    return np.abs(a.astype(np.float64) - ref.astype(np.float64)).max() / (np.finfo(np.float32).eps * scale)


def _call(lib, fn, x):
    from pointcloudprocessing_b200 import _ffi
    b, n, _ = x.shape
    y = torch.full_like(x, float("nan"))
    c = torch.full((b, 3), float("nan"), device=x.device)
    s = torch.full((b,), float("nan"), device=x.device)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _ffi.check(fn(_ffi.ptr(x), b, n, _ffi.ptr(y), _ffi.ptr(c), _ffi.ptr(s), _ffi.stream_ptr()))
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return y.cpu().numpy(), c.cpu().numpy(), s.cpu().numpy(), free0 - free1


def test_normalize_noalloc_matches_existing_entry_point_twice(lib):
    from pointcloudprocessing_b200 import _ffi
    assert lib.pcp_normalize_noalloc_max_clouds() == 4096
    taken = []
    for seed, b, n in ((31, 6, 40000), (32, 3, 16384)):
        xh = synth.gaussian_clouds(seed, b, n)
        x = torch.from_numpy(xh).cuda()
        y0, c0, s0, _ = _call(lib, lib.pcp_normalize_f32, x)
        y1, c1, s1, used = _call(lib, lib.pcp_normalize_noalloc_f32, x)
        taken.append(used)
        assert np.isfinite(y1).all() and np.isfinite(c1).all() and np.isfinite(s1).all()
        # the tolerances of tests/test_gpu_preprocess.py::test_normalize_matches_oracle (same kernel: normally bit-identical)
        assert _ulp_err(c1, c0, np.abs(c0).max() + 1e-30) <= 1.0
        assert _ulp_err(s1, s0, np.abs(s0).max()) <= 2.0
        assert _ulp_err(y1, y0, 1.0) <= max(4.0 * (np.abs(xh).max() / s0.min()), 4.0)
    # the second call (kernel module loaded, per-device set-up done) must not have taken any device memory
    assert taken[1] == 0, taken


def test_normalize_noalloc_rejects_what_the_static_scratch_cannot_hold(lib):
    from pointcloudprocessing_b200 import _ffi
    x = torch.zeros((1, 16384, 3), device="cuda")
    y = torch.empty_like(x)
    args = (_ffi.ptr(y), None, None, _ffi.stream_ptr())
    with pytest.raises(_ffi.PcpError, match="exceeds the static scratch"):
        _ffi.check(lib.pcp_normalize_noalloc_f32(_ffi.ptr(x), 4097, 16384, *args))
    with pytest.raises(_ffi.PcpError, match="only the sliced kernel uses scratch"):
        _ffi.check(lib.pcp_normalize_noalloc_f32(_ffi.ptr(x), 1, 1000, *args))
    assert lib.pcp_normalize_noalloc_f32(_ffi.ptr(x), 0, 16384, *args) == 0
