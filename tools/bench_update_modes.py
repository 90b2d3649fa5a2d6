"""K4 micro-benchmark of one tile mode at one size (for ncu captures): python tools/bench_update_modes.py [mode] [rows]"""
import ctypes as C, sys
sys.path.insert(0, ".")
from uno_b200 import capi
lib = capi.load_library()
lib.b200kkt_debug_bench_update.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
mode = int(sys.argv[1]) if len(sys.argv) > 1 else 4
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
s, fl = C.c_double(), C.c_double()
rc = lib.b200kkt_debug_bench_update(0, rows + 256, 256, 3, mode, C.byref(s), C.byref(fl))
print(f"mode {mode} rows {rows} K=256: rc={rc} {s.value*1e3:.3f} ms {fl.value/s.value/1e12:.2f} TFLOP/s, algorithmic flops {fl.value:.4e}")
