"""Development: per-k-tile slope of the persistent GEMM with loads and stores compiled out (dbg=3)."""
import ctypes as C, sys
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import _ffi
lib = _ffi.load()
rows = 262144
for k, n in ((64, 128), (128, 128), (256, 128), (512, 128), (1024, 128), (64, 64), (256, 64), (1024, 64), (64, 32), (1024, 32)):
    for dbg in (3, 0):
        ms = C.c_float()
        _ffi.check(lib.pcp_bench_tc_gemm(rows, k, n, 1, 0, 1, 10, dbg, C.byref(ms)))
        tiles = rows / 128 / 148
        print(f"dbg={dbg} K={k:4d} N={n:4d}: {ms.value * 1e3:8.1f} us  = {ms.value * 1e3 / tiles:6.2f} us/tile", flush=True)
