"""The CPU oracle against the reference's golden vectors and dense numpy (CPU only)."""
import json
from pathlib import Path

import numpy as np
import pytest

from helpers import backward_error, dense_inertia, kat_system
from oracle import oracle as O
from uno_b200 import generators as G

GOLDEN = Path(__file__).resolve().parent / "golden"
KATS = json.loads((GOLDEN / "reference_kats.json").read_text())
VECTORS = json.loads((GOLDEN / "oracle_vectors.json").read_text())


def oracle_run(kkt, values=None, perm=None, **kw):
    s = O.OracleSolver(**kw)
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, perm)
    s.do_numerical_factorization(kkt.val if values is None else values)
    return s


def test_kat1_solve_reference_vector():
    # unotest/functional_tests/MUMPSSolverTests.cpp:11-38 (tolerance 1e-8 there)
    e = KATS["kat1_solve"]
    kkt = kat_system(e)
    x = oracle_run(kkt).solve_indefinite_system(kkt.rhs)
    assert np.max(np.abs(x - np.array(e["x"]))) <= e["tol"]


def test_kat2_inertia_reference_vector():
    # MUMPSSolverTests.cpp:40-61: inertia (3, 2, 0); eigenvalues in the comment at :86-100
    e = KATS["kat2_inertia"]
    kkt = kat_system(e)
    assert list(oracle_run(kkt).get_inertia()) == e["inertia"]
    w = np.linalg.eigvalsh(kkt.to_scipy().toarray())
    assert np.allclose(np.sort(w), np.sort(e["eigenvalues"]), atol=1e-6)


def test_kat3_singular_with_duplicates_reference_vector():
    # MUMPSSolverTests.cpp:63-83: duplicates are summed, matrix_is_singular(), expected inertia (1, 1, 2)
    e = KATS["kat3_singular"]
    s = oracle_run(kat_system(e))
    assert s.matrix_is_singular()
    assert list(s.get_inertia()) == e["inertia"]


@pytest.mark.parametrize("name", sorted(VECTORS))
def test_oracle_matches_committed_vectors(name):
    mk = {
        "ocp2d_k12": lambda: G.ocp_distributed(12, 2, seed=1),
        "ocp3d_k6": lambda: G.ocp_distributed(6, 3, seed=2),
        "ocp3d_boundary_k6": lambda: G.ocp_boundary_3d(6, seed=3),
        "opf_like_b40": lambda: G.opf_like(40, 55, 8, seed=4),
        "opf_like_b40_reg": lambda: G.opf_like(40, 55, 8, seed=4, delta_w=50.0),
    }[name]
    kkt, e = mk(), VECTORS[name]
    assert (kkt.n, kkt.m, kkt.nnz) == (e["n"], e["m"], e["nnz"])
    s = oracle_run(kkt)
    assert list(s.get_inertia()) == e["inertia"]
    x = s.solve_indefinite_system(kkt.rhs)
    assert np.allclose(x[:8], e["x_head"], rtol=1e-7, atol=1e-9 * e["x_norm_inf"])
    assert backward_error(kkt, kkt.val, x) < 1e-13


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_oracle_inertia_against_dense_eigenvalues(seed):
    kkt = G.opf_like(50, 70, 10, seed=seed)
    for dw in (0.0, 1.0, 100.0):
        v = kkt.with_regularization(dw, 1e-8 if dw else 0.0)
        assert oracle_run(kkt, v).get_inertia() == dense_inertia(kkt.to_scipy(v).toarray())


def test_oracle_ordering_independence_and_threads():
    kkt = G.ocp_distributed(10, 2, seed=5)
    a = oracle_run(kkt)
    rng = np.random.default_rng(0)
    b = oracle_run(kkt, perm=rng.permutation(kkt.dim), threads=4)
    assert a.get_inertia() == b.get_inertia() == (kkt.n, kkt.m, 0)
    xa, xb = a.solve_indefinite_system(kkt.rhs), b.solve_indefinite_system(kkt.rhs)
    assert backward_error(kkt, kkt.val, xa) < 1e-13 and backward_error(kkt, kkt.val, xb) < 1e-13


def test_oracle_rank_deficient_jacobian_reports_zero_pivots():
    kkt = G.opf_like(40, 55, 8, seed=7, duplicate_rows=3)
    v = kkt.with_regularization(100.0, 0.0)
    s = oracle_run(kkt, v, null_tol=1e-12)
    pos, neg, zero = s.get_inertia()
    assert zero == 3 and pos + neg + zero == kkt.dim
    # with dual regularization the matrix is nonsingular again and the inertia is (n, m, 0)
    v = kkt.with_regularization(100.0, 1e-6)
    assert oracle_run(kkt, v, null_tol=1e-12).get_inertia() == (kkt.n, kkt.m, 0)


def _inertia_of_block_diagonal(d):
    """Inertia of LAPACK's block-diagonal D (1x1 and 2x2 blocks) from the eigenvalues of its blocks."""
    n = d.shape[0]
    pos = neg = zero = 0
    i = 0
    while i < n:
        if i + 1 < n and d[i + 1, i] != 0.0:
            w = np.linalg.eigvalsh(d[i:i + 2, i:i + 2])
            i += 2
        else:
            w = np.array([d[i, i]])
            i += 1
        pos += int((w > 0).sum()); neg += int((w < 0).sum()); zero += int((w == 0).sum())
    return pos, neg, zero


@pytest.mark.parametrize("restricted", [False, True])
@pytest.mark.parametrize("name", ["ocp2d_26", "opf_400", "bnd3d_9"])
def test_oracle_against_lapack_bunch_kaufman(name, restricted):
    """An anchor outside this repository: LAPACK's dsytrf (Bunch-Kaufman partial pivoting, the dense kernel of the algorithm family
    MUMPS belongs to) through scipy.linalg.ldl on KKT systems of 1500 - 2100 rows.  Same inertia (Sylvester: it is a property of the
    matrix, so any stable LDL^T must agree) and the same solution to rounding, in both pivoting modes of the oracle."""
    import scipy.linalg as sla
    kkt = {"ocp2d_26": lambda: G.ocp_distributed(26, 2, seed=3), "opf_400": lambda: G.opf_like(400, 540, 80, seed=4, delta_w=10.0),
           "bnd3d_9": lambda: G.ocp_boundary_3d(9, seed=5)}[name]()
    # the OPF-like matrix gets the dual regularization the interior-point loop would give it (without it rcond ~ 1e-17 and the sign
    # of one eigenvalue is rounding noise: LAPACK and the oracle's restricted mode then differ by one)
    v = kkt.with_regularization(10.0, 1e-8) if name == "opf_400" else kkt.val
    a = kkt.to_scipy(v).toarray()
    _, d, _ = sla.ldl(a, lower=True)
    s = oracle_run(kkt, v, restricted=restricted)
    assert tuple(s.get_inertia()) == _inertia_of_block_diagonal(d) == (kkt.n, kkt.m, 0)
    x = s.solve_indefinite_system(kkt.rhs)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")          # the OPF-like matrix without dual regularization has rcond ~ 1e-17
        x_lapack = sla.solve(a, kkt.rhs, assume_a="sym")
    assert backward_error(kkt, v, x) < 1e-13
    # two backward-stable solutions differ by about cond * eps; the barrier diagonals of these systems span 1e-8 .. 1e8
    assert np.max(np.abs(x - x_lapack)) <= 1e-6 * np.max(np.abs(x_lapack))


def test_oracle_against_superlu_at_baseline_config_1():
    """Second outside anchor, at a BASELINE size: SuperLU (scipy.sparse.linalg.splu, an unrelated sparse LU with its own ordering and
    partial pivoting) on the 256 x 256 control problem, N = 196 608: same solution to rounding times the condition number."""
    import scipy.sparse.linalg as spla
    kkt = G.ocp_distributed(256, 2, seed=0)
    lu = spla.splu(kkt.to_scipy().tocsc())
    x_lu = lu.solve(kkt.rhs)
    assert backward_error(kkt, kkt.val, x_lu) < 1e-10
    # MUMPS' rule (threshold test against the whole front column, 225 000 delayed pivots here): backward error at rounding level.
    # Restricted mode (no delays) on the oracle's own ordering, no refinement: 2e-11 - the accuracy that the product's value-aware
    # pairing and refinement are there to recover.
    for restricted, berr_max, diff_max in ((False, 1e-15, 1e-7), (True, 1e-9, 1e-4)):
        s = oracle_run(kkt, restricted=restricted, threads=4)
        assert tuple(s.get_inertia()) == (kkt.n, kkt.m, 0)
        x = s.solve_indefinite_system(kkt.rhs)
        assert backward_error(kkt, kkt.val, x) < berr_max
        assert np.max(np.abs(x - x_lu)) <= diff_max * np.max(np.abs(x_lu))


@pytest.mark.parametrize("restricted", [False, True])
@pytest.mark.parametrize("n,m,seed", [(60, 40, 0), (200, 120, 1), (500, 499, 2)])
def test_oracle_null_pivot_count_on_a_saddle_point_matrix_without_hessian(n, m, seed, restricted):
    """INFOG(28) (MUMPSSolver.cpp:138) on a matrix whose rank deficiency is known in closed form: [[0, A^T], [A, 0]] with A of full
    row rank m has m positive, m negative and n - m zero eigenvalues.  Every pivot of such a matrix is a 2x2 block or a null pivot,
    which exercises the parts of the pivoting the regularized KKT systems rarely reach; the dense spectrum is the referee."""
    import scipy.sparse as sp
    a = (sp.random(m, n, density=min(0.15, 8.0 / n), random_state=seed, format="coo") + sp.eye(m, n)).tocoo()
    row = list(range(n + m)) + [int(j) for j in a.col]
    col = list(range(n + m)) + [int(n + i) for i in a.row]
    val = [0.0] * (n + m) + [float(v) for v in a.data]
    row, col, val = np.array(row, dtype=np.uint64), np.array(col, dtype=np.uint64), np.array(val)
    k = sp.coo_matrix((val, (row.astype(np.int64), col.astype(np.int64))), shape=(n + m, n + m))
    w = np.linalg.eigvalsh((k + sp.triu(k, 1).T).toarray())
    tol = 1e-9 * np.abs(w).max()
    dense = (int((w > tol).sum()), int((w < -tol).sum()), int((np.abs(w) <= tol).sum()))
    assert dense == (m, m, n - m)
    s = O.OracleSolver(restricted=restricted, null_tol=1e-12)
    s.do_symbolic_analysis(n + m, row, col, None)
    s.do_numerical_factorization(val)
    assert tuple(s.get_inertia()) == dense
