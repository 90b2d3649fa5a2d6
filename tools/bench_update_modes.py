import ctypes as C, sys
sys.path.insert(0, ".")
from uno_b200 import capi
lib = capi.load_library()
lib.b200kkt_debug_bench_update.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
for mode in (3, 4):
    s, fl = C.c_double(), C.c_double()
    rc = lib.b200kkt_debug_bench_update(0, 8192 + 256, 256, 3, mode, C.byref(s), C.byref(fl))
    print(mode, rc, s.value * 1e3, fl.value / s.value / 1e12)
