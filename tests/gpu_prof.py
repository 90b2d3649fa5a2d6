"""One un-graphed factorization + solve of one workload, for ncu launch lists (run under gpurun)."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from uno_b200 import KKTSolver
from uno_b200 import generators as G

name = sys.argv[1] if len(sys.argv) > 1 else "k32"
kkt = {"k32": lambda: G.ocp_distributed(32, 3), "k48": lambda: G.ocp_distributed(48, 3), "k64": lambda: G.ocp_distributed(64, 3),
       "k256": lambda: G.ocp_distributed(256, 2), "opf": lambda: G.opf_like(delta_w=100.0)}[name]()
s = KKTSolver(device=0, use_graph=0)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
s.do_numerical_factorization(kkt.val)
x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
st = s.stats()
print(f"{kkt.name}: factor {st['t_factor_last']*1e3:.1f} ms solve {st['t_solve_last']*1e3:.1f} ms inertia={s.get_inertia()} berr={berr:.2e} launches={st['kernel_launches']}")
