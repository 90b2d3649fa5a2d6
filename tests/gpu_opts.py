"""Factor time of one workload under option variants, graph mode (run under gpurun): python tests/gpu_opts.py k64 lookahead=0 lookahead=1"""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from uno_b200 import KKTSolver
from uno_b200 import generators as G
name = sys.argv[1]
kkt = {"k32": lambda: G.ocp_distributed(32, 3), "k48": lambda: G.ocp_distributed(48, 3), "k64": lambda: G.ocp_distributed(64, 3),
       "k256": lambda: G.ocp_distributed(256, 2), "opf": lambda: G.opf_like(delta_w=100.0)}[name]()
for variant in sys.argv[2:] or [""]:
    s = KKTSolver(device=0)
    for kv in variant.split(","):
        if kv:
            k, v = kv.split("=")
            s.set_option(k, float(v))
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    ts = []
    for _ in range(4):
        s.do_numerical_factorization(kkt.val)
        ts.append(s.stats()["t_factor_last"] * 1e3)
    tsol = []
    for _ in range(3):
        x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
        tsol.append(s.stats()['t_solve_last'] * 1e3)
    print(f"{kkt.name} [{variant}]: factor {min(ts[1:]):.1f} ms (first {ts[0]:.1f}) solve {min(tsol[1:]):.1f} ms (first {tsol[0]:.1f}) inertia={s.get_inertia()} berr={berr:.2e}", flush=True)
    del s
