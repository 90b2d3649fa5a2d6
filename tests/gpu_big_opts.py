"""128^3 boundary-control system: factor time under option variants on one analysis (run under gpurun):
python tests/gpu_big_opts.py 128 pair_panels=1 pair_panels=0"""
import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from uno_b200 import KKTSolver
from uno_b200 import generators as G
k = int(sys.argv[1])
kkt = G.ocp_boundary_3d(k)
s = KKTSolver(device=0)
t0 = time.time()
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
print(f"analysed {kkt.name} in {time.time()-t0:.0f} s", flush=True)
for variant in sys.argv[2:]:
    for kv in variant.split(","):
        a, b = kv.split("=")
        s.set_option(a, float(b))
    ts = []
    for _ in range(2):
        s.do_numerical_factorization(kkt.val)
        ts.append(s.stats()["t_factor_last"])
    st = s.stats()
    print(f"[{variant}] factor {min(ts):.3f} s (first {ts[0]:.3f}) {st['flops_factor']/min(ts)/1e12:.2f} TFLOP/s inertia={s.get_inertia()}", flush=True)
x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
print(f"solve {s.stats()['t_solve_last']:.3f} s berr={berr:.2e}")
