"""Phase cycle counters of the panel kernels (needs the -DB200KKT_PHASES build variant; run under gpurun)."""
import ctypes as C, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from uno_b200 import KKTSolver, capi
from uno_b200 import generators as G
name = sys.argv[1] if len(sys.argv) > 1 else "k48"
kkt = {"k32": lambda: G.ocp_distributed(32, 3), "k48": lambda: G.ocp_distributed(48, 3), "k64": lambda: G.ocp_distributed(64, 3)}[name]()
s = KKTSolver(device=0, use_graph=0)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
s.do_numerical_factorization(kkt.val)
lib = capi.load_library()
out = (C.c_ulonglong * 16)()
lib.b200kkt_debug_phase_cycles(out, 1)
s.do_numerical_factorization(kkt.val)
rc = lib.b200kkt_debug_phase_cycles(out, 0)
tot = sum(out[:8])
names = ["diag load", "bunch-kaufman", "writeback+interchange", "row trsm", "in-block update", "", "", "", "bk: loop overhead", "bk: pivot search", "bk: interchange", "bk: 1x1 update", "bk: 2x2 update"]
print("compiled in:", rc, "factor ms", s.stats()["t_factor_last"] * 1e3)
for i, n in enumerate(names):
    print(f"  {n:24s} {out[i]/1e6:10.2f} Mcycles {100*out[i]/max(tot,1):5.1f}%")
