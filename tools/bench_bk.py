"""Cycles per 32x32 Bunch-Kaufman block (run under gpurun)."""
import ctypes as C, sys
sys.path.insert(0, ".")
from uno_b200 import capi
lib = capi.load_library()
for v, name in enumerate(["dominant diagonal", "KKT-like (tiny every 2nd diagonal)", "random symmetric"]):
    c = C.c_longlong()
    rc = lib.b200kkt_debug_bench_bk(v, 200, C.byref(c))
    print(f"{name:40s} rc={rc} {c.value} cycles / block  ({c.value/32:.0f} per column)")
