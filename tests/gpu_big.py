"""BASELINE configs[4] on ONE GPU: 3D boundary-control KKT, 128^3 grid (N = 4.29 M).  Run under gpurun; several minutes of host analysis."""
import faulthandler, sys, time
faulthandler.dump_traceback_later(1500, repeat=False)
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from uno_b200 import KKTSolver
from uno_b200 import generators as G

k = int(sys.argv[1]) if len(sys.argv) > 1 else 128
t0 = time.time()
kkt = G.ocp_boundary_3d(k)
print(f"[{time.time()-t0:7.1f}s] generated {kkt.name}: N={kkt.dim} n={kkt.n} m={kkt.m} nnz={kkt.nnz}", flush=True)
s = KKTSolver(device=0)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
st = s.stats(); p = s.plan()
print(f"[{time.time()-t0:7.1f}s] analysed: fronts={p['num_fronts']} levels={p['num_levels']} factor pool={p['pool_doubles']*8/1e9:.1f} GB "
      f"CB arena={p['cb_arena_doubles']*8/1e9:.1f} GB nnz(L)={st['nnz_factor']:.3e} flops={st['flops_factor']:.3e} max front={st['max_front']} "
      f"(ordering {st['t_ordering']:.0f}s symbolic {st['t_symbolic']:.0f}s maps {st['t_maps']:.0f}s upload {st['t_upload']:.0f}s)", flush=True)
try:
    import torch
    free, total = torch.cuda.mem_get_info(0)
    print(f"device memory: {(total-free)/1e9:.1f} GB in use of {total/1e9:.1f} GB", flush=True)
except Exception as exc:
    print("mem_get_info unavailable:", exc)
for it in range(2):
    s.do_numerical_factorization(kkt.val)
    st = s.stats()
    print(f"[{time.time()-t0:7.1f}s] factorize {it}: {st['t_factor_last']:.3f} s  {st['flops_factor']/st['t_factor_last']/1e12:.2f} TFLOP/s inertia={s.get_inertia()} expected={(kkt.n, kkt.m, 0)}", flush=True)
for it in range(2):
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    st = s.stats()
    print(f"[{time.time()-t0:7.1f}s] solve {it}: {st['t_solve_last']:.3f} s refine={st['refine_steps_last']} berr={berr:.2e}", flush=True)
r = kkt.rhs - kkt.matvec(x) if hasattr(kkt, "matvec") else None
if r is not None:
    print("host residual inf-norm / rhs inf-norm:", np.max(np.abs(r)) / np.max(np.abs(kkt.rhs)))
