"""Development: one tc_gemm shape for ncu.  args: rows k n kind relu dbg iters"""
import ctypes as C, sys
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import _ffi
lib = _ffi.load()
rows, k, n, kind, relu, dbg, iters = [int(a) for a in sys.argv[1:8]]
ms = C.c_float()
_ffi.check(lib.pcp_bench_tc_gemm(rows, k, n, kind, 0, relu, iters, dbg, C.byref(ms)))
print(f"K={k} N={n} kind={kind} dbg={dbg}: {ms.value * 1e3:.1f} us")
