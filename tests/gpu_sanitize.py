"""One pass over the numeric path on small systems, meant to run under compute-sanitizer on the GPU box:

  compute-sanitizer --tool memcheck --error-exitcode 9 python tests/gpu_sanitize.py

Covers K1 (fused and scatter), K2, K3a/K3b (both row-panel kernels), K4, K5, K6, the small- and big-front solves (with and without
gather lists), refinement, the diagonal-shift fast path, a factorization with delayed pivots (re-plan) and the device-side
interior-point steps.  Results are checked as in the parity suite (inertia, backward error); sizes keep the run to a few minutes
under the tool."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import torch  # noqa: E402

from uno_b200 import KKTSolver  # noqa: E402
from uno_b200 import generators as G  # noqa: E402


def run(kkt, expect=None, **options):
    s = KKTSolver(device=0, **options)
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
    s.do_numerical_factorization(kkt.val)
    ine = tuple(s.get_inertia())
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    if kkt.reg_size > 0:
        s.refactorize_shifted(kkt.val[:kkt.reg_size])
        assert tuple(s.get_inertia()) == ine
    st = s.stats()
    print(f"{kkt.name} {options}: N={kkt.dim} inertia={ine} berr={berr:.1e} max_front={st['max_front']} delayed={st['num_delayed_total']}", flush=True)
    assert expect is None or ine == tuple(expect), (ine, expect)
    assert berr <= 1e-12 and np.all(np.isfinite(x))
    s.close()


def main():
    k = G.ocp_distributed(16, 3, seed=0)                      # top fronts beyond one outer panel: K3a/K3b/K4 panel loop, big-front solves
    run(k, (k.n, k.m, 0))
    run(k, (k.n, k.m, 0), use_graph=0, gather_lists=0, fused_assembly=0, fused_leaves=0, lookahead=0)
    k = G.ocp_boundary_3d(14, seed=1)
    run(k, (k.n, k.m, 0))
    k = G.ocp_distributed(40, 2, seed=2)
    run(k, (k.n, k.m, 0))
    k = G.opf_like(300, 405, 60, seed=3, delta_w=100.0)       # delayed pivots: flags, host re-plan, second attempt
    run(k, (k.n, k.m, 0))
    k = G.opf_like(300, 405, 60, seed=3, delta_w=100.0)
    run(k, (k.n, k.m, 0), positive_definite_target=0, growth_limit=1e6)
    # device-side interior-point steps on a resident system
    k = G.ocp_distributed(12, 2, seed=4)
    s = KKTSolver(device=0)
    s.do_symbolic_analysis(k.dim, k.row, k.col, k.reg_size, k.n, values=k.val)
    s.do_numerical_factorization(k.val)
    n, m = k.n, k.m
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()   # noqa: E731
    g, c, y = dev(np.ones(n)), dev(np.ones(m)), dev(np.linspace(-1, 1, m))
    rhs = torch.empty(n + m, dtype=torch.float64, device="cuda")
    s.assemble_rhs_device(g.data_ptr(), c.data_ptr(), y.data_ptr(), rhs.data_ptr())
    prod = torch.empty_like(rhs)
    s.symmetric_product_device(rhs.data_ptr(), prod.data_ptr())
    x = torch.empty_like(rhs)
    s.solve_device(rhs.data_ptr(), x.data_ptr())
    s.shift_diagonal_device(n, 1e-4, 1e-8)
    s.factorize_resident()
    assert tuple(s.get_inertia()) == (n, m, 0)
    nw, cw = s.backward_error_noalloc(x.data_ptr(), rhs.data_ptr())
    assert np.isfinite(nw) and np.isfinite(cw)
    torch.cuda.synchronize()
    s.close()
    print("sanitize pass: all checks green", flush=True)


if __name__ == "__main__":
    main()
