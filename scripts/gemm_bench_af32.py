import ctypes as C, sys
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import _ffi
lib = _ffi.load()
ms = C.c_float()
rows, k, n, kind, a_f32, dbg = [int(a) for a in sys.argv[1:7]]
_ffi.check(lib.pcp_bench_tc_gemm(rows, k, n, kind, a_f32, 0, 5, dbg, C.byref(ms)))
print(ms.value * 1e3)
