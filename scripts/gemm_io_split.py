import ctypes as C, sys
sys.path.insert(0, ".")
from pointcloudprocessing_b200 import _ffi
lib = _ffi.load()
ms = C.c_float()
for (k, n) in ((64, 64), (64, 128), (128, 64), (128, 128)):
    for a_f32 in (0, 1):
        for kind in (0, 1):
            row = []
            for dbg in (0, 1, 2, 3):
                _ffi.check(lib.pcp_bench_tc_gemm(262144, k, n, kind, a_f32, 0, 20, dbg, C.byref(ms)))
                row.append(ms.value * 1e3)
            mb = 262144 * (k * (4 if a_f32 else 2) + n * (2 if kind else 4)) / 1e6
            print(f"K={k:3d} N={n:3d} A={'f32' if a_f32 else 'bf16'} C={'bf16' if kind else 'f32 '}: {row[0]:6.1f} us ({mb / row[0] * 1e-6 * 1e6:5.2f} TB/s) | no stores {row[1]:6.1f} | no loads {row[2]:6.1f} | neither {row[3]:6.1f}")
