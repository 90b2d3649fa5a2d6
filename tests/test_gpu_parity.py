"""GPU parity tests: the CUDA path through the C ABI against the CPU oracle, the reference's golden
vectors and size-independent properties.  Integer results (assembly permutation, inertia triple,
pivot sequence) must be bit-exact; the solution must reach a relative backward error <= 1e-12
(BASELINE.json north_star)."""
import json
from pathlib import Path

import numpy as np
import pytest

from emulator import EmulatedFactorization, equilibrate
from helpers import backward_error, inertia_correction_loop, kat_system
from oracle import oracle as O
from uno_b200 import B200KKTError, KKTSolver
from uno_b200 import generators as G

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"
KATS = json.loads((GOLDEN / "reference_kats.json").read_text())
BERR_TOL = 1e-12    # north_star: relative backward error after refinement
X_TOL = 1e-4        # agreement with the oracle's solution (the reference's own CI tolerance: runtests_uno_ipopt_mumps.jl:72-73)


def gpu_solver(kkt, **opts):
    s = KKTSolver(device=0, **opts)
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n if kkt.m else 0, values=kkt.val)
    return s


def oracle_solver(kkt, values, perm=None, **kw):
    o = O.OracleSolver(**kw)
    o.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, perm)
    o.do_numerical_factorization(values)
    return o


def test_kat1_solution():
    e = KATS["kat1_solve"]
    kkt = kat_system(e)
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    assert np.max(np.abs(s.solve_indefinite_system(kkt.rhs) - np.array(e["x"]))) <= e["tol"]


def test_kat2_inertia():
    e = KATS["kat2_inertia"]
    kkt = kat_system(e)
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    assert list(s.get_inertia()) == e["inertia"]
    assert s.number_negative_eigenvalues() == 2 and not s.matrix_is_singular() and s.rank() == 5


def test_kat3_singular_duplicates():
    e = KATS["kat3_singular"]
    kkt = kat_system(e)
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    assert s.matrix_is_singular() and list(s.get_inertia()) == e["inertia"] and s.rank() == 2


CASES = {
    "ocp2d_k24": lambda: G.ocp_distributed(24, 2, seed=1),
    "ocp3d_k9": lambda: G.ocp_distributed(9, 3, seed=2),
    "ocp3d_boundary_k9": lambda: G.ocp_boundary_3d(9, seed=3),
    "opf_like_b150": lambda: G.opf_like(150, 200, 30, seed=4),
    "opf_like_b150_reg": lambda: G.opf_like(150, 200, 30, seed=4, delta_w=100.0),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_assembly_permutation_bit_exact(name):
    kkt = CASES[name]()
    s = gpu_solver(kkt)
    perm, scat = s.permutation()
    pool, scale = s.debug_assemble(kkt.val)
    assert np.array_equal(scale, equilibrate(kkt, kkt.val, 3))
    ref = np.zeros_like(pool)
    r, c = kkt.row.astype(np.int64), kkt.col.astype(np.int64)
    np.add.at(ref, scat, kkt.val * scale[r] * scale[c])     # duplicates summed in ascending slot order
    assert np.array_equal(pool, ref)
    # and the destinations really are P K P^T: rebuild the matrix from the pool and compare with scipy
    plan = s.plan()
    iperm = np.empty(kkt.dim, dtype=np.int64)
    iperm[perm] = np.arange(kkt.dim)
    dense = np.zeros((kkt.dim, kkt.dim))
    for sn in range(plan["num_fronts"]):
        k, f, ld, off = plan["k"][sn], plan["f"][sn], plan["ld"][sn], plan["pool_offset"][sn]
        rows = plan["rows"][plan["rows_off"][sn] : plan["rows_off"][sn + 1]]
        F = pool[off : off + ld * k].reshape(k, ld).T[:f, :k]
        for j in range(k):
            dense[rows[j:], rows[j]] += F[j:, j]
    dense = dense + np.tril(dense, -1).T
    A = kkt.to_scipy(kkt.val).toarray() * scale[:, None] * scale[None, :]
    # scipy sums duplicate COO entries in its own order: allow one rounding of the largest addend per entry
    absA = (abs(kkt.to_scipy(np.abs(kkt.val))).toarray() * scale[:, None] * scale[None, :])[np.ix_(perm, perm)]
    assert np.all(np.abs(dense - A[np.ix_(perm, perm)]) <= 4e-16 * absA)


@pytest.mark.parametrize("name", sorted(CASES))
def test_inertia_and_solution_against_oracle(name):
    kkt = CASES[name]()
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    perm, _ = s.permutation()
    for o in (oracle_solver(kkt, kkt.val), oracle_solver(kkt, kkt.val, perm=perm)):     # own METIS ordering, and the GPU's ordering
        assert tuple(s.get_inertia()) == tuple(o.get_inertia())
        x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
        assert berr <= BERR_TOL and backward_error(kkt, kkt.val, x) <= BERR_TOL
        xo = o.solve_indefinite_system(kkt.rhs)
        assert np.max(np.abs(x - xo)) <= X_TOL * np.max(np.abs(xo))


def test_pivot_sequence_matches_plan_emulation():
    """Same plan, same pivoting rule -> identical pivot types and in-front permutations (integers: bit-exact)."""
    kkt = G.opf_like(120, 160, 20, seed=9)
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    perm, scat = s.permutation()
    emu = EmulatedFactorization(s.plan(), perm, scat, kkt, kkt.val)
    _, dd, de, pt, lp = s.debug_factor(want_pool=False)
    assert np.array_equal(pt, emu.ptype) and np.array_equal(lp, emu.lperm)
    assert tuple(s.get_inertia()) == emu.inertia()


def test_refactorize_shifted_equals_full_factorize_bitwise():
    kkt = G.opf_like(150, 200, 30, seed=5)
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.with_regularization(0.0, 0.0))
    v = kkt.with_regularization(3.7, 1e-8)
    s.refactorize_shifted(v[: kkt.reg_size])
    a = s.debug_factor()
    xa = s.solve_indefinite_system(kkt.rhs)
    ia = s.get_inertia()
    s2 = gpu_solver(kkt, use_graph=0)
    s2.do_numerical_factorization(v)
    b = s2.debug_factor()
    assert ia == s2.get_inertia()
    for u, w in zip(a[1:], b[1:]):
        assert np.array_equal(u, w)
    assert np.array_equal(xa, s2.solve_indefinite_system(kkt.rhs))     # graph replay == plain launches, bit for bit


def test_factorization_is_deterministic():
    kkt = G.ocp_distributed(12, 3, seed=6)
    s = gpu_solver(kkt, adaptive_growth=0)      # the plan must not adapt between the two factorizations
    s.do_numerical_factorization(kkt.val)
    a = s.debug_factor()
    xa = s.solve_indefinite_system(kkt.rhs)
    s.do_numerical_factorization(kkt.val)
    b = s.debug_factor()
    assert all(np.array_equal(u, w) for u, w in zip(a, b))
    assert np.array_equal(xa, s.solve_indefinite_system(kkt.rhs))


def test_inertia_correction_loop_matches_oracle_sequence():
    """Indefinite Hessian: the caller's loop must see the same verdicts from both solvers, hence the same delta sequence."""
    kkt = G.opf_like(150, 200, 30, seed=8, barrier_exp=(-8.0, 0.0))
    s = gpu_solver(kkt)
    hist_gpu, v_gpu = inertia_correction_loop(s, kkt)
    o = O.OracleSolver()
    o.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col)
    hist_cpu, _ = inertia_correction_loop(o, kkt)
    assert len(hist_gpu) > 1 and hist_gpu == hist_cpu
    assert s.get_inertia() == (kkt.n, kkt.m, 0)
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    assert berr <= BERR_TOL
    # the loop only changed the regularization slots: the fast path gives the same verdicts
    s2 = gpu_solver(kkt)
    state = {"first": True}

    def fact(values):
        if state["first"]:
            s2.do_numerical_factorization(values)
            state["first"] = False
        else:
            s2.refactorize_shifted(values[: kkt.reg_size])

    hist_fast, _ = inertia_correction_loop(s2, kkt, factorize=fact)
    assert hist_fast == hist_gpu


def test_rank_deficient_jacobian_triggers_dual_regularization():
    kkt = G.opf_like(120, 160, 20, seed=7, duplicate_rows=4)
    s = gpu_solver(kkt)
    # a duplicated row leaves a pivot of rounding size, not an exact zero: both solvers get the same threshold for it
    s.set_option("null_pivot_tol", 1e-12)
    s.do_numerical_factorization(kkt.with_regularization(100.0, 0.0))
    pos, neg, zero = s.get_inertia()
    o = oracle_solver(kkt, kkt.with_regularization(100.0, 0.0), null_tol=1e-12)
    assert o.get_inertia() == (kkt.n, kkt.m - 4, 4)
    # What the caller branches on (PrimalDualRegularization.hpp:157,164): "inertia wrong" and "singular" - both must agree with
    # the oracle.  The COUNT of zero pivots can come out lower than the nullity: the pivot search is confined to the
    # fully-summed block, the elimination of two identical rows then takes different routes and their last pivot is rounding
    # noise times the element growth instead of an exact zero (known limits).
    assert s.matrix_is_singular() and 1 <= zero <= 4 and pos + neg + zero == kkt.dim and s.rank() == kkt.dim - zero
    assert (pos, neg, zero) != (kkt.n, kkt.m, 0)
    hist, _ = inertia_correction_loop(s, kkt)
    o2 = O.OracleSolver(null_tol=1e-12)
    o2.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col)
    hist_cpu, _ = inertia_correction_loop(o2, kkt)
    assert hist == hist_cpu and hist[-1][1] > 0.0 and s.get_inertia() == (kkt.n, kkt.m, 0)


def test_two_interleaved_handles_and_reanalysis():
    # optimality and feasibility IPMs each own a solver (FeasibilityRestoration.cpp:33-34); the pre-solve
    # re-analyses another pattern on the same instance (Preprocessing.cpp:60)
    a, b = G.ocp_distributed(10, 2, seed=1), G.opf_like(60, 80, 12, seed=2, delta_w=100.0)
    sa, sb = gpu_solver(a), gpu_solver(b)
    sa.do_numerical_factorization(a.val)
    sb.do_numerical_factorization(b.val)
    xa = sa.solve_indefinite_system(a.rhs)
    xb = sb.solve_indefinite_system(b.rhs)
    assert backward_error(a, a.val, xa) <= BERR_TOL and backward_error(b, b.val, xb) <= BERR_TOL
    sub = slice(a.reg_size, None)
    a2 = type(a)("noreg", a.n, a.m, 0, a.row[sub], a.col[sub], a.val[sub], a.rhs)
    sa.do_symbolic_analysis(a2.dim, a2.row, a2.col, 0, a2.n)
    with pytest.raises(B200KKTError):
        sa.solve_indefinite_system(a2.rhs)          # new pattern: the old factors are gone
    sa.do_numerical_factorization(a2.val)
    assert sa.get_inertia() == (a.n, a.m, 0)
    assert backward_error(a2, a2.val, sa.solve_indefinite_system(a2.rhs)) <= BERR_TOL
    assert backward_error(b, b.val, sb.solve_indefinite_system(b.rhs)) <= BERR_TOL   # the other handle is untouched


def test_hessian_only_positive_definite_path():
    # byrd preset: PrimalRegularization factorizes the Hessian alone (m = 0) and expects (n, 0, 0)
    rng = np.random.default_rng(3)
    n = 300
    import scipy.sparse as sp
    M = sp.random(n, n, density=0.02, random_state=4, format="coo")
    M = sp.triu(M + M.T, 1).tocoo()
    row = np.concatenate([np.arange(n), M.row]).astype(np.uint64)
    col = np.concatenate([np.arange(n), M.col]).astype(np.uint64)
    val = np.concatenate([np.zeros(n), M.data])
    kkt = G.KKTSystem("hess", n, 0, n, row, col, val, rng.standard_normal(n))
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    assert s.get_inertia() != (n, 0, 0)
    shift = 1.1 * abs(M).sum(axis=1).max() + abs(M).sum(axis=0).max()
    s.refactorize_shifted(np.full(n, shift))
    assert s.get_inertia() == (n, 0, 0)
    v = kkt.with_regularization(shift, 0.0)
    assert backward_error(kkt, v, s.solve_indefinite_system(kkt.rhs)) <= BERR_TOL


def test_config3_delayed_pivots_give_the_exact_inertia():
    """BASELINE configs[2] at full size (10k buses, delta_w = 100, delta_c = 0).  Three multiplier columns have no pivot inside
    the fully-summed block of their front (J restricted to the front is rank deficient): the factorization flags them, the host
    moves them into an ancestor front and factorizes again.  Structural inertia (n, m, 0); round 1 reported (51000, 46997, 3)."""
    kkt = G.opf_like(delta_w=100.0)
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    st = s.stats()
    assert s.get_inertia() == (kkt.n, kkt.m, 0) == (51000, 47000, 0)
    assert st["num_delayed_total"] >= 1 and st["num_replans_total"] >= 1 and st["num_null_pivots_last"] == 0
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    assert berr <= BERR_TOL and backward_error(kkt, kkt.val, x) <= BERR_TOL
    assert s.stats()["backward_error_componentwise_last"] <= BERR_TOL
    # the moves persist: the next factorization of the same pattern starts from the adapted plan
    s.do_numerical_factorization(kkt.val)
    assert s.stats()["num_replans_last"] == 0 and s.get_inertia() == (kkt.n, kkt.m, 0)
    assert np.array_equal(x, s.solve_indefinite_system(kkt.rhs))
    # and so does the inertia-loop fast path
    s.refactorize_shifted(kkt.val[: kkt.reg_size])
    assert s.get_inertia() == (kkt.n, kkt.m, 0)


BASELINE_CASES = {
    "config2_ocp2d_k256": lambda: G.ocp_distributed(256, 2),
    "config3_opf_10k": lambda: G.opf_like(delta_w=100.0),
    "config4_ocp3d_k64": lambda: G.ocp_distributed(64, 3),
}


@pytest.mark.parametrize("name", sorted(BASELINE_CASES))
def test_baseline_sizes_against_the_live_oracle(name):
    """BASELINE configs[1..3] at FULL size, GPU against the CPU oracle run live on the same ordering:
    inertia triple equal (and equal to the structural (n, m, 0)), both backward errors <= 1e-12, solutions agree.  The oracle runs
    in its restricted-pivoting mode (threshold test over the fully-summed rows, delayed pivots kept): its full-column mode needs
    the better part of an hour at these sizes."""
    import os
    kkt = BASELINE_CASES[name]()
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    st = s.stats()
    perm, _ = s.permutation()
    O.bind_openblas()
    o = oracle_solver(kkt, kkt.val, perm=perm, threads=os.cpu_count() or 1, restricted=True)
    assert tuple(s.get_inertia()) == tuple(o.get_inertia()) == (kkt.n, kkt.m, 0)
    assert berr <= BERR_TOL and st["backward_error_componentwise_last"] <= BERR_TOL and backward_error(kkt, kkt.val, x) <= BERR_TOL
    assert st["refine_steps_last"] <= 2
    xo = o.solve_indefinite_system(kkt.rhs)
    assert np.max(np.abs(x - xo)) <= X_TOL * np.max(np.abs(xo))


def test_config3_with_duplicated_rows_gives_the_singular_verdict():
    """configs[2], 10 constraint rows duplicated (rank-deficient Jacobian), delta_c = 0: the caller must be told "inertia wrong"
    and "singular" (PrimalDualRegularization.hpp:157,164-167), as the oracle tells it.  Under restricted pivoting neither the
    GPU nor the CPU oracle finds all ten zero pivots (see test_rank_deficient_jacobian_triggers_dual_regularization)."""
    import os
    kkt = G.opf_like(delta_w=100.0, duplicate_rows=10)
    s = gpu_solver(kkt, null_pivot_tol=1e-12)
    s.do_numerical_factorization(kkt.val)
    pos, neg, zero = s.get_inertia()
    perm, _ = s.permutation()
    O.bind_openblas()
    o = oracle_solver(kkt, kkt.val, perm=perm, threads=os.cpu_count() or 1, restricted=True, null_tol=1e-12)
    assert s.matrix_is_singular() and o.matrix_is_singular()
    assert 1 <= zero <= 10 and pos + neg + zero == kkt.dim
    assert (pos, neg, zero) != (kkt.n, kkt.m, 0) != tuple(o.get_inertia())
    # with the dual regularization the caller then applies (delta_c = 1e-8 mu^0.25 at mu = 1) both are regular with the right inertia
    v = kkt.with_regularization(100.0, 1e-8)
    s.do_numerical_factorization(v)
    o.do_numerical_factorization(v)
    assert tuple(s.get_inertia()) == tuple(o.get_inertia()) == (kkt.n, kkt.m, 0)


def test_non_finite_values_are_an_error_not_an_inertia():
    """MUMPS would return INFOG(1) < 0; a NaN pivot must never be counted as a negative eigenvalue."""
    kkt = G.ocp_distributed(8, 2, seed=5)
    s = gpu_solver(kkt)
    bad = kkt.val.copy()
    bad[kkt.reg_size + 3] = np.nan
    with pytest.raises(B200KKTError):
        s.do_numerical_factorization(bad)
    s.do_numerical_factorization(kkt.val)          # the handle stays usable
    assert s.get_inertia() == (kkt.n, kkt.m, 0)
    rhs = kkt.rhs.copy()
    rhs[0] = np.inf
    with pytest.raises(B200KKTError):
        s.solve_indefinite_system(rhs)


def test_positive_definite_target_skips_the_pivot_search():
    """Cholesky-type LDL^T (no search, no interchanges, 1 x 1 pivots) for the Hessian-convexification path.
    Same verdicts as the pivoting factorization along the caller's shift sequence (PrimalRegularization.hpp:90-129), same
    solution once the matrix is positive definite."""
    import scipy.sparse as sp
    k = 48
    n = k * k
    t = sp.diags([-np.ones(k - 1), np.zeros(k), -np.ones(k - 1)], [-1, 0, 1])
    H = sp.triu(sp.kron(sp.identity(k), t) + sp.kron(t, sp.identity(k)) + 2.5 * sp.identity(n)).tocoo()     # 5-point stencil, diagonal 2.5: indefinite
    row = np.concatenate([np.arange(n), H.row]).astype(np.uint64)
    col = np.concatenate([np.arange(n), H.col]).astype(np.uint64)
    kkt = G.KKTSystem("hessian", n, 0, n, row, col, np.concatenate([np.zeros(n), H.data]), np.random.default_rng(1).standard_normal(n))
    ref, pd = gpu_solver(kkt), gpu_solver(kkt, positive_definite_target=1)
    delta, verdicts = 0.0, []
    while True:
        v = kkt.with_regularization(delta, 0.0)
        ref.do_numerical_factorization(v)
        pd.do_numerical_factorization(v)
        ok_ref, ok_pd = ref.get_inertia() == (n, 0, 0), pd.get_inertia() == (n, 0, 0)
        assert ok_ref == ok_pd
        verdicts.append(ok_pd)
        if ok_pd:
            break
        delta = 1e-4 if delta == 0.0 else 2.0 * delta
    assert len(verdicts) > 5 and delta > 1.0
    _, _, _, ptype, lperm = pd.debug_factor(want_pool=False)
    first = pd.plan()["first_col"]
    assert np.all(ptype == 0)                                                   # 1 x 1 pivots only ...
    assert np.array_equal(lperm, np.arange(n) - np.repeat(first, pd.plan()["k"]))   # ... in the given order
    x, berr = pd.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    xr = ref.solve_indefinite_system(kkt.rhs)
    assert berr <= BERR_TOL and np.max(np.abs(x - xr)) <= 1e-10 * np.max(np.abs(xr))


@pytest.mark.parametrize("make,name", [(lambda: G.ocp_distributed(256, 2), "config2"), (lambda: G.ocp_distributed(32, 3), "3d_k32")])
def test_full_size_properties(make, name):
    """At BASELINE sizes the oracle is too slow: check structural inertia (convex OCP: exactly (n, m, 0)),
    backward error, linearity of the solve and refactorization idempotence."""
    kkt = make()
    s = gpu_solver(kkt)
    s.do_numerical_factorization(kkt.val)
    assert s.get_inertia() == (kkt.n, kkt.m, 0)
    x1, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    assert berr <= BERR_TOL
    rng = np.random.default_rng(0)
    b2 = rng.standard_normal(kkt.dim)
    x2 = s.solve_indefinite_system(b2)
    x12 = s.solve_indefinite_system(kkt.rhs + 2.0 * b2)
    assert backward_error(kkt, kkt.val, x12, kkt.rhs + 2.0 * b2) <= BERR_TOL
    # forward error ~ cond(K) * backward error; the barrier terms spread over 16 decades, so only the reference's own 1e-4 is asked
    assert np.max(np.abs(x12 - (x1 + 2.0 * x2))) <= X_TOL * np.max(np.abs(x12))
    # idempotence once the plan has settled (the first solve may arm the growth guard: the next factorization then moves the
    # columns responsible for the growth up the tree, and the plan stays as it is from there on)
    s.do_numerical_factorization(kkt.val)
    x3, berr3 = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    assert s.get_inertia() == (kkt.n, kkt.m, 0) and berr3 <= BERR_TOL
    assert s.stats()["backward_error_componentwise_last"] <= BERR_TOL
    if name == "config2":
        assert s.stats()["refine_steps_last"] <= 2      # config 2 needed 6 sweeps
    s.do_numerical_factorization(kkt.val)
    assert s.stats()["num_replans_last"] == 0
    assert np.array_equal(x3, s.solve_indefinite_system(kkt.rhs))


@pytest.mark.gpu
@pytest.mark.parametrize("options", [
    {"lookahead": 0, "multi_stream": 0, "pair_panels": 0, "use_graph": 0},       # plain level-by-level, one stream
    {"lookahead": 1, "pair_panels": 0},                                           # look-ahead by single panels
    {"lookahead": 1, "pair_panels": 1, "pair_min_front": 512},                    # two outer panels per trailing update
    {"lookahead": 0, "pair_panels": 1, "pair_min_front": 512},
    {"update_tile_mode": 1}, {"update_tile_mode": 2}, {"update_tile_mode": 3}, {"update_tile_mode": 6},
    {"gather_lists": 0},                                                          # forward gather by the children loop instead of the precomputed lists
    {"small_solve_ctas": 0},                                                      # 256-thread CTAs on every level of the small-front sweeps
    {"fused_assembly": 0},                                                        # memset + scatter instead of the fused K1
    {"fused_leaves": 0}, {"fused_leaves": 1},                                     # small fronts through K3a/K3b/K4 instead of K2
])
def test_schedule_and_kernel_variants_agree(options):
    """Streams, look-ahead, panel pairing and the K4 tile variants only reorder independent work (and, for pairing, the
    summation order of the trailing update): same pivots, same inertia, same solution to rounding."""
    kkt = G.ocp_distributed(26, 3, seed=3)         # fronts up to ~2000 rows, several outer panels per top front
    ref = gpu_solver(kkt)
    ref.do_numerical_factorization(kkt.val)
    xr, berr_r = ref.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    s = gpu_solver(kkt)
    for name, value in options.items():
        s.set_option(name, value)
    s.do_numerical_factorization(kkt.val)
    x, berr = s.solve_indefinite_system(kkt.rhs, return_backward_error=True)
    assert s.get_inertia() == ref.get_inertia() == (kkt.n, kkt.m, 0)
    assert berr <= BERR_TOL and berr_r <= BERR_TOL
    if "pair_min_front" not in options and "update_tile_mode" not in options and "fused_leaves" not in options:
        # pure re-scheduling: every entry of the factor sees the same operations in the same order
        a, b = ref.debug_factor(), s.debug_factor()
        assert all(np.array_equal(u, w) for u, w in zip(a, b))
        assert np.array_equal(x, xr)
    assert np.max(np.abs(x - xr)) <= X_TOL * np.max(np.abs(xr))


@pytest.mark.gpu
def test_partitioned_two_gpus_matches_single_gpu():
    """the same system on two GPUs (subtrees mapped to the ranks, top fronts shared, NCCL over NVLink) gives the
    inertia and, to rounding, the solution of the one-GPU run; both ranks return the same x bit for bit.  Needs two devices:
    one process per GPU under torchrun (tests/gpu_partition.py does the comparison and exits non-zero on a mismatch)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices (run with gpurun --gpus 2)")
    root = Path(__file__).resolve().parent.parent
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1", "--master-port", "29533",
           str(root / "tests" / "gpu_partition.py"), "ocp3d:16", "bnd3d:20", "ocp3d:32", "share_min_pivots=512", "share_min_flops=1e9"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    lines = [json.loads(l) for l in r.stdout.replace("}{", "}\n{").splitlines() if l.startswith("{")]
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert len(lines) == 6 and all(l["ok"] for l in lines)
    assert any(l["shared_fronts"] > 0 for l in lines)        # the shared-front path ran, not only the subtree exchange
    assert all(l["pool_MB"] < l["pool_MB_whole_system"] for l in lines)      # every rank allocates its own part of the factors only
