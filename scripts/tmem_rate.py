"""Development: TMEM read rate of one CTA (tcgen05.ld 32x32b.x32, four loads per wait)."""
import ctypes as C, sys
sys.path.insert(0, ".")
import numpy as np
from pointcloudprocessing_b200 import _ffi
lib = C.CDLL(str(_ffi.lib_path()))
lib.pcp_selftest_tmem_ld_rate.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
for nw in (1, 2, 4):
    for pack in (0, 1):
        out = np.zeros(8, np.int64)
        assert lib.pcp_selftest_tmem_ld_rate(2000, nw, pack, out.ctypes.data_as(C.c_void_p)) == 0
        print("warps", nw, "pack16", pack, "clk/iter per warp", out[:4] / 2000.0, flush=True)
