"""Development: time the generic tensor-core GEMM alone on the segmentation-head layer shapes (C4: 128 x 2048 points)."""
import ctypes as C
import sys
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import _ffi

lib = _ffi.load()
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
shapes = [(64, 512, 1, 1), (512, 256, 1, 1), (256, 128, 1, 1), (128, 128, 1, 1), (128, 12, 2, 0)]
for dbg in (0, 1, 2, 3):
    for k, n, kind, relu in shapes:
        ms = C.c_float()
        _ffi.check(lib.pcp_bench_tc_gemm(rows, k, n, kind, 0, relu, 20, dbg, C.byref(ms)))
        byts = rows * (k * 2 + n * (2 if kind == 1 else 4))
        print(f"dbg={dbg} K={k:4d} N={n:4d} kind={kind}: {ms.value * 1e3:8.1f} us  {byts / ms.value / 1e9:7.2f} TB/s (algorithmic)", flush=True)
