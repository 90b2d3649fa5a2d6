"""K4 trailing-update microbenchmark through the C ABI hook (run under gpurun)."""
import ctypes as C, sys
sys.path.insert(0, ".")
from uno_b200 import capi
import os
lib = C.CDLL(os.environ['B200KKT_LIB']) if os.environ.get('B200KKT_LIB') else capi.load_library()
lib.b200kkt_debug_bench_update.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
sizes = [int(a) for a in sys.argv[1:]] or [2048, 4096, 8192, 12288]
for f in sizes:
    for k in (256, 64):
        for mode in (1, 2, 3, 4, 5):
            s, fl = C.c_double(), C.c_double()
            rc = lib.b200kkt_debug_bench_update(0, f + k, k, 5, mode, C.byref(s), C.byref(fl))
            print(f"update r={f} K={k} tiles={['', '128x128', '128x64x2', 'ws', '128x64x2 8 warps', '128x128 16 warps'][mode]}: rc={rc} {s.value*1e3:.3f} ms  {fl.value/s.value/1e12:.2f} TFLOP/s (lower triangle only)", flush=True)
