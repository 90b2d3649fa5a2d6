"""b200kkt_backward_error_noalloc: the allocation-free entry point (static device scratch) against the existing one.

Collected last on purpose (a failure here hides nothing).  The existing entry point
b200kkt_solve_device reports the normwise and componentwise backward error of the x it returns; the new one must give the same
two numbers for that x, on two calls in a row (the scratch array keeps the previous call's contents), and again after another
system went through the same array."""
import numpy as np
import pytest

from uno_b200 import B200KKTError, KKTSolver
from uno_b200 import generators as G

pytestmark = pytest.mark.gpu
TOL = 1e-12     # the tolerance of the GPU parity tests (tests/test_gpu_parity.py: BERR_TOL)


def _solve_on_device(torch, kkt):
    s = KKTSolver(device=0)
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
    s.do_numerical_factorization(kkt.val)
    d_b = torch.from_numpy(kkt.rhs).cuda()
    d_x = torch.empty_like(d_b)
    berr = s.solve_device(d_b.data_ptr(), d_x.data_ptr())
    omega = s.stats()["backward_error_componentwise_last"]
    torch.cuda.synchronize()
    return s, d_b, d_x, berr, omega


def test_no_alloc_backward_error_matches_solve_twice_in_a_row():
    import torch
    kkt = G.ocp_distributed(20, 2, seed=11)
    s, d_b, d_x, berr, omega = _solve_on_device(torch, kkt)
    assert berr <= TOL and omega <= TOL
    for _ in range(2):
        nw, cw = s.backward_error_noalloc(d_x.data_ptr(), d_b.data_ptr())
        assert abs(nw - berr) <= TOL and abs(cw - omega) <= TOL
        assert nw == berr and cw == omega          # same kernels on the same data: in fact identical
    # a perturbed x: both measures must see it, and the stale contents of the scratch array must not leak into the next call
    d_bad = d_x.clone()
    d_bad[3] += 1.0
    nw_bad, cw_bad = s.backward_error_noalloc(d_bad.data_ptr(), d_b.data_ptr())
    assert nw_bad > 1e3 * max(berr, 1e-20) and cw_bad > 1e3 * max(omega, 1e-20)
    # another (smaller) system through the same static array, then the first one again
    kkt2 = G.ocp_distributed(6, 3, seed=12)
    s2, d_b2, d_x2, berr2, omega2 = _solve_on_device(torch, kkt2)
    assert s2.backward_error_noalloc(d_x2.data_ptr(), d_b2.data_ptr()) == (berr2, omega2)
    assert s.backward_error_noalloc(d_x.data_ptr(), d_b.data_ptr()) == (berr, omega)
    # host check of the normwise number
    x = d_x.cpu().numpy()
    A = kkt.to_scipy()
    r = kkt.rhs - A @ x
    ref = np.abs(r).max() / (abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(kkt.rhs).max())
    assert abs(ref - berr) <= TOL


def test_no_alloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    n = 1048576 + 1                       # B200KKT_NOALLOC_MAX_N + 1: a diagonal system
    idx = np.arange(n, dtype=np.uint64)
    s = KKTSolver(device=0)
    s.do_symbolic_analysis(n, idx, idx, 0, 0)
    s.do_numerical_factorization(np.full(n, 2.0))
    d = torch.ones(n, dtype=torch.float64, device="cuda")
    with pytest.raises(B200KKTError, match="B200KKT_NOALLOC_MAX_N"):
        s.backward_error_noalloc(d.data_ptr(), d.data_ptr())
