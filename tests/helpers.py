import numpy as np


def backward_error(kkt, values, x, rhs=None):
    rhs = kkt.rhs if rhs is None else rhs
    A = kkt.to_scipy(values)
    r = rhs - A @ x
    return np.abs(r).max() / (abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(rhs).max())


def dense_inertia(A, tol=1e-11):
    """Inertia from eigenvalues of the equilibrated matrix (congruence keeps the inertia)."""
    A = np.array(A, dtype=np.float64)
    for _ in range(4):
        m = np.abs(A).max(axis=1)
        d = np.where(m > 0, 1.0 / np.sqrt(np.where(m > 0, m, 1.0)), 1.0)
        A = A * d[:, None] * d[None, :]
    w = np.linalg.eigvalsh(A)
    return int((w > tol).sum()), int((w < -tol).sum()), int((np.abs(w) <= tol).sum())


def kat_system(entry):
    from uno_b200.generators import KKTSystem
    n = entry["n"]
    return KKTSystem("kat", n, 0, 0, np.array(entry["row"], dtype=np.uint64), np.array(entry["col"], dtype=np.uint64),
                     np.array(entry["val"], dtype=np.float64), np.array(entry.get("rhs", [1.0] * n), dtype=np.float64))


def inertia_correction_loop(solver, kkt, dual_regularization_parameter=1.0, previous_primal=0.0, factorize=None):
    """Restatement of the caller's inertia-correction loop
    (uno/ingredients/regularization_strategies/PrimalDualRegularization.hpp:133-219, default options
    uno/options/DefaultOptions.cpp:109-122).  Returns the (delta_w, delta_c) sequence that was
    factorized and the final values array.  `factorize(values)` defaults to
    solver.do_numerical_factorization."""
    init, lb, dec, fast, slow = 1e-4, 1e-20, 3.0, 100.0, 8.0
    dual_fraction, threshold_attempts, failure = 1e-8, 8, 1e40
    factorize = factorize or solver.do_numerical_factorization
    expected = (kkt.n, kkt.m, 0)
    dw = dc = 0.0
    attempts = 1
    values = kkt.with_regularization(0.0, 0.0)
    history = [(dw, dc)]
    factorize(values)
    if tuple(solver.get_inertia()) == expected:
        return history, values
    if solver.matrix_is_singular():
        dc = dual_fraction * dual_regularization_parameter
    dw = init if previous_primal == 0.0 else max(lb, previous_primal / dec)
    while True:
        values = kkt.with_regularization(dw, dc)
        history.append((dw, dc))
        factorize(values)
        attempts += 1
        if tuple(solver.get_inertia()) == expected:
            return history, values
        dw *= fast if (previous_primal == 0.0 or threshold_attempts < attempts) else slow
        if dw > failure:
            raise RuntimeError("UnstableRegularization")
