"""Development: tcgen05.mma issue rate (cycles per MMA) for one CTA and for a CTA pair."""
import ctypes as C
from pathlib import Path
lib = C.CDLL(str(Path(__file__).resolve().parents[1] / "pointcloudprocessing_b200/lib/libpcp_b200.so"))
out = (C.c_longlong * 10)()
for N in (64, 128, 256):
    lib.pcp_selftest_umma_rate(N, 4000, 2, 1, 1, out)
    ld = [(out[2 + 2 * w], out[3 + 2 * w]) for w in range(4)]
    bpc = sum(n * 16384 / c for c, n in ld if c)
    print(f"1-CTA  M=128 N={N}: {out[0]/out[1]:.1f} cycles/MMA (floor {128*N/256:.0f}); concurrent tcgen05.ld {bpc:.0f} B/clk/SM")
for ncl in (1, 74):
    for N in (64, 128, 256):
        lib.pcp_selftest_umma2_rate(N, 4000, 2, ncl, out)
        print(f"pair x{ncl} M=256 N={N}: {out[0]/out[1]:.1f} cycles/MMA (floor {256*N/512:.0f})")
