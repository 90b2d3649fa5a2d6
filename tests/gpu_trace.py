"""Ad-hoc GPU timing trace of one workload (run under gpurun)."""
import faulthandler, sys, time
faulthandler.dump_traceback_later(90, repeat=True)
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from uno_b200 import KKTSolver
from uno_b200 import generators as G

def T(msg, t0):
    print(f"[{time.time()-t0:8.2f}s] {msg}", flush=True)

name = sys.argv[1] if len(sys.argv) > 1 else "k48"
graph = int(sys.argv[2]) if len(sys.argv) > 2 else 1
t0 = time.time()
kkt = {"k32": lambda: G.ocp_distributed(32, 3), "k48": lambda: G.ocp_distributed(48, 3), "k64": lambda: G.ocp_distributed(64, 3),
       "k256": lambda: G.ocp_distributed(256, 2), "opf": lambda: G.opf_like(delta_w=100.0), "b96": lambda: G.ocp_boundary_3d(96)}[name]()
T("generated " + kkt.name, t0)
s = KKTSolver(device=0, use_graph=graph)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
st = s.stats()
T(f"analysed fronts={st['num_fronts']} maxf={st['max_front']} pool={st['pool_doubles']*8/1e9:.1f}GB flops={st['flops_factor']:.3g}", t0)
for i in range(3):
    s.do_numerical_factorization(kkt.val)
    st = s.stats()
    T(f"factorize {i}: dev {st['t_factor_last']*1e3:.1f} ms  {st['flops_factor']/st['t_factor_last']/1e12:.2f} TF/s inertia={s.get_inertia()} launches={st['kernel_launches']}", t0)
for i in range(2):
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    st = s.stats()
    T(f"solve {i}: dev {st['t_solve_last']*1e3:.1f} ms refine={st['refine_steps_last']} berr={berr:.2e}", t0)
