"""Regenerates tests/golden/*.json.  Two kinds of vectors:
 * reference KATs: copied by hand from the reference's dormant solver tests
   (unotest/functional_tests/MUMPSSolverTests.cpp:11-83, MA27SolverTests.cpp:11-36) — the only
   results the reference pins on this path;
 * oracle vectors: inertia / solution of seeded synthetic KKT systems computed with the CPU oracle
   (oracle/kkt_oracle.cpp) AND cross-checked against dense numpy (eigvalsh / solve) at generation
   time.  They guard the oracle against regressions; the GPU tests compare against the live oracle.
Run from the repo root:  python tests/golden/make_golden.py
"""
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle import oracle as O  # noqa: E402
from uno_b200 import generators as G  # noqa: E402

OUT = Path(__file__).resolve().parent


def dense_inertia(A, tol=1e-11):
    """Inertia from dense eigenvalues of the symmetrically equilibrated matrix (a congruence, so
    Sylvester's law keeps the inertia); the raw KKT matrices span 16 decades."""
    A = np.array(A, dtype=np.float64)
    for _ in range(4):
        m = np.abs(A).max(axis=1)
        d = np.where(m > 0, 1.0 / np.sqrt(np.where(m > 0, m, 1.0)), 1.0)
        A = A * d[:, None] * d[None, :]
    w = np.linalg.eigvalsh(A)
    return [int((w > tol).sum()), int((w < -tol).sum()), int((np.abs(w) <= tol).sum())]


def main():
    kats = {
        "kat1_solve": {"source": "unotest/functional_tests/MUMPSSolverTests.cpp:11-38", "n": 5,
                       "row": [0, 0, 1, 1, 2, 2, 4], "col": [0, 1, 2, 4, 2, 3, 4], "val": [2.0, 3.0, 4.0, 6.0, 1.0, 5.0, 1.0],
                       "rhs": [8.0, 45.0, 31.0, 15.0, 17.0], "x": [1.0, 2.0, 3.0, 4.0, 5.0], "tol": 1e-8},
        "kat2_inertia": {"source": "unotest/functional_tests/MUMPSSolverTests.cpp:40-61,86-100", "n": 5,
                         "row": [0, 0, 1, 1, 2, 2, 4], "col": [0, 1, 2, 4, 2, 3, 4], "val": [2.0, 3.0, 4.0, 6.0, 1.0, 5.0, 1.0],
                         "inertia": [3, 2, 0], "eigenvalues": [-7.83039207, 8.94059148, -3.50815575, 1.7888887, 4.60906763]},
        "kat3_singular": {"source": "unotest/functional_tests/MUMPSSolverTests.cpp:63-83", "n": 4,
                          "row": [0, 0, 0, 1, 1, 2, 3], "col": [0, 0, 1, 1, 1, 2, 3],
                          "val": [-0.0198, 0.625075, -0.277512, -0.624975, 0.625075, 0.0, 0.0], "inertia": [1, 1, 2], "singular": True},
    }
    (OUT / "reference_kats.json").write_text(json.dumps(kats, indent=1))

    cases = {
        "ocp2d_k12": lambda: G.ocp_distributed(12, 2, seed=1),
        "ocp3d_k6": lambda: G.ocp_distributed(6, 3, seed=2),
        "ocp3d_boundary_k6": lambda: G.ocp_boundary_3d(6, seed=3),
        "opf_like_b40": lambda: G.opf_like(40, 55, 8, seed=4),
        "opf_like_b40_reg": lambda: G.opf_like(40, 55, 8, seed=4, delta_w=50.0),
    }
    out = {}
    for name, mk in cases.items():
        kkt = mk()
        s = O.OracleSolver()
        s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col)
        s.do_numerical_factorization(kkt.val)
        x = s.solve_indefinite_system(kkt.rhs)
        A = kkt.to_scipy().toarray()
        ine = dense_inertia(A)
        assert list(s.get_inertia()) == ine, (name, s.get_inertia(), ine)
        xd = np.linalg.solve(A, kkt.rhs)
        assert np.allclose(x, xd, rtol=1e-6, atol=1e-9 * np.abs(xd).max()), name
        out[name] = {"n": kkt.n, "m": kkt.m, "nnz": kkt.nnz, "inertia": ine, "x_head": x[:8].tolist(), "x_norm_inf": float(np.abs(x).max()),
                     "values_checksum": float(np.sum(kkt.val * np.arange(1, kkt.nnz + 1) % 7))}
    (OUT / "oracle_vectors.json").write_text(json.dumps(out, indent=1))
    print("wrote", sorted(p.name for p in OUT.glob("*.json")))


if __name__ == "__main__":
    main()
