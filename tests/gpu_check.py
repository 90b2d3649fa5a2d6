"""Ad-hoc GPU sanity script (run under gpurun): GPU solver vs numpy emulator / scipy on small and medium systems."""
import sys, time
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from uno_b200 import KKTSolver
from uno_b200 import generators as G
from emulator import EmulatedFactorization


def check(kkt, vals=None, emu=True, graph=1, expect=None):
    vals = kkt.val if vals is None else vals
    s = KKTSolver(device=0, use_graph=graph)
    t0 = time.time()
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n if kkt.m else 0)
    t1 = time.time()
    s.do_numerical_factorization(vals)
    t2 = time.time()
    s.do_numerical_factorization(vals)
    t3 = time.time()
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    t4 = time.time()
    A = kkt.to_scipy(vals)
    r = kkt.rhs - A @ x
    berr_host = np.abs(r).max() / (abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(kkt.rhs).max())
    st = s.stats()
    msg = f"{kkt.name}: N={kkt.dim} inertia={s.get_inertia()} berr={berr:.2e} host_berr={berr_host:.2e} refine={st['refine_steps_last']} " \
          f"analyse={t1-t0:.2f}s fact1={t2-t1:.3f}s fact2={t3-t2:.4f}s (dev {st['t_factor_last']*1e3:.2f}ms) solve={t4-t3:.4f}s (dev {st['t_solve_last']*1e3:.2f}ms) " \
          f"fronts={st['num_fronts']} maxf={st['max_front']} flops={st['flops_factor']:.3g} GF/s={st['flops_factor']/max(st['t_factor_last'],1e-9)/1e9:.1f} launches={st['kernel_launches']}"
    if emu:
        perm, scat = s.permutation()
        E = EmulatedFactorization(s.plan(), perm, scat, kkt, vals)
        pool, scale = s.debug_assemble(vals)
        msg += f" | assembly_bitexact={np.array_equal(pool, E.assembled)} scale_equal={np.array_equal(scale, E.scale)} emu_inertia={E.inertia()}"
        s.do_numerical_factorization(vals)
        _, dd, de, pt, lp = s.debug_factor(want_pool=False)
        msg += f" ptype_equal={np.array_equal(pt, E.ptype)} lperm_equal={np.array_equal(lp, E.lperm)} dd_relerr={np.max(np.abs(dd-E.dd)/np.maximum(np.abs(E.dd),1e-300)):.1e}"
        xe = E.solve(kkt.rhs)
        msg += f" x_vs_emu={np.max(np.abs(x-xe))/max(np.max(np.abs(xe)),1e-300):.1e}"
    if expect is not None:
        msg += f" expect={expect} {'OK' if tuple(s.get_inertia()) == tuple(expect) else 'MISMATCH'}"
    print(msg, flush=True)
    s.close()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "small"
    if which in ("small", "all"):
        k = G.kat_matrix(); check(k, expect=(3, 2, 0), graph=0)
        k = G.kat_singular(); check(k, expect=(1, 1, 2))
        k = G.ocp_distributed(16, 2); check(k, expect=(k.n, k.m, 0)); check(k, k.with_regularization(1e-4, 1e-8), expect=(k.n, k.m, 0))
        k = G.ocp_distributed(8, 3); check(k, expect=(k.n, k.m, 0))
        k = G.ocp_boundary_3d(8); check(k, expect=(k.n, k.m, 0))
        k = G.opf_like(60, 80, 12); check(k); check(k, k.with_regularization(1e2, 1e-8))
        k = G.ocp_distributed(48, 2); check(k, expect=(k.n, k.m, 0))
        k = G.ocp_distributed(14, 3); check(k, expect=(k.n, k.m, 0))
    if which in ("medium", "all"):
        k = G.ocp_distributed(256, 2); check(k, emu=False, expect=(k.n, k.m, 0))
        k = G.ocp_distributed(32, 3); check(k, emu=False, expect=(k.n, k.m, 0))
        k = G.opf_like(); check(k, k.with_regularization(1e2, 1e-8), emu=False)
    if which in ("large", "all"):
        k = G.ocp_distributed(64, 3); check(k, emu=False, expect=(k.n, k.m, 0))
