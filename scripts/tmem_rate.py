import ctypes as C, sys
from pathlib import Path
lib = C.CDLL(str(Path(__file__).resolve().parents[1] / "pointcloudprocessing_b200/lib/libpcp_b200.so"))
out = (C.c_longlong * 8)()
for pack in (0, 1):
    for nw in (1, 2, 4):
        lib.pcp_selftest_tmem_ld_rate(1000, nw, pack, out)
        cyc = max(out[i] for i in range(nw))
        per_warp_bytes = 1000 * (4096 if pack else 16384)
        print(f"pack16={pack} warps={nw}: {cyc/1000:.1f} cycles per iteration per warp; {per_warp_bytes*nw/cyc:.1f} B/clk per SM")
