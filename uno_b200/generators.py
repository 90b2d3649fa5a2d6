"""Deterministic synthetic KKT systems for the BASELINE.json configs.

Every generator returns the augmented matrix exactly as Uno hands it to the linear solver:
upper-triangular COO, entry order = [regularization slots (k,k) | Lagrangian Hessian, column by
column | barrier diagonals | Jacobian entries (var, n+j), constraint by constraint]
(uno/linear_algebra/COOFormat.hpp:119-125, uno/ingredients/subproblem/Subproblem.cpp:57-70,
uno/ingredients/inequality_handling_methods/interior_point_methods/PrimalDualInteriorPointProblem.cpp:56-78).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import scipy.sparse as sp


@dataclass
class KKTSystem:
    name: str
    n: int            # primal variables
    m: int            # equality constraints
    reg_size: int     # leading regularization slots
    row: np.ndarray   # uint64, row <= col
    col: np.ndarray   # uint64
    val: np.ndarray   # float64, analysis order
    rhs: np.ndarray   # float64, length n + m

    @property
    def dim(self) -> int:
        return self.n + self.m

    @property
    def nnz(self) -> int:
        return int(self.row.size)

    def with_regularization(self, delta_w: float, delta_c: float) -> np.ndarray:
        """Values after set_regularization(primal, 0, +dw) / (dual, n, -dc)
        (uno/ingredients/regularization_strategies/PrimalDualRegularization.hpp:178-179)."""
        v = self.val.copy()
        if self.reg_size >= self.n + self.m:
            v[: self.n] = delta_w
            v[self.n : self.n + self.m] = -delta_c
        elif self.reg_size > 0:
            v[: self.reg_size] = delta_w
        return v

    def to_scipy(self, values=None) -> sp.csr_matrix:
        """Full symmetric matrix (duplicates summed)."""
        v = self.val if values is None else values
        r = self.row.astype(np.int64)
        c = self.col.astype(np.int64)
        off = r != c
        rr = np.concatenate([r, c[off]])
        cc = np.concatenate([c, r[off]])
        vv = np.concatenate([v, v[off]])
        return sp.coo_matrix((vv, (rr, cc)), shape=(self.dim, self.dim)).tocsr()


def assemble_kkt(name, n, m, hess_row, hess_col, hess_val, barrier_idx, barrier_val, jac: sp.csr_matrix, rhs, reg_size=None) -> KKTSystem:
    """Lay the pieces out in the reference's COO order."""
    reg = n + m if reg_size is None else reg_size
    jac = jac.tocsr()
    jac.sort_indices()
    jrow = np.repeat(np.arange(m, dtype=np.int64), np.diff(jac.indptr))
    row = np.concatenate([np.arange(reg, dtype=np.int64), hess_row, barrier_idx, jac.indices.astype(np.int64)])
    col = np.concatenate([np.arange(reg, dtype=np.int64), hess_col, barrier_idx, n + jrow])
    val = np.concatenate([np.zeros(reg), hess_val, barrier_val, jac.data.astype(np.float64)])
    assert np.all(row <= col)
    return KKTSystem(name, int(n), int(m), int(reg), row.astype(np.uint64), col.astype(np.uint64), val.astype(np.float64),
                     np.asarray(rhs, dtype=np.float64))


def _laplacian(k: int, dim: int) -> sp.csr_matrix:
    """-Laplace_h with homogeneous Dirichlet conditions on a k^dim grid, scaled by h^-2."""
    h2inv = float((k + 1) ** 2)
    t = sp.diags([-np.ones(k - 1), 2.0 * np.ones(k), -np.ones(k - 1)], [-1, 0, 1], format="csr")
    eye = sp.identity(k, format="csr")
    if dim == 2:
        lap = sp.kron(eye, t) + sp.kron(t, eye)
    else:
        lap = sp.kron(sp.kron(eye, eye), t) + sp.kron(sp.kron(eye, t), eye) + sp.kron(sp.kron(t, eye), eye)
    return (h2inv * lap).tocsr()


def _barrier_sigma(rng, count, lo=-8.0, hi=8.0):
    # late-IPM spread of z/(x-l): log-uniform over 16 decades by default
    return 10.0 ** rng.uniform(lo, hi, size=count)


def ocp_distributed(k: int, dim: int, alpha: float = 1e-3, seed: int = 0) -> KKTSystem:
    """min 1/2|y-yd|^2 + alpha/2|u|^2  s.t.  -Lap_h y - u = f,  u in [-1,1]   (configs 2 and 4).
    x = [y ; u], one equality per grid node."""
    rng = np.random.default_rng(seed)
    nodes = k ** dim
    n, m = 2 * nodes, nodes
    hess_idx = np.arange(n, dtype=np.int64)
    hess_val = np.concatenate([np.ones(nodes), alpha * np.ones(nodes)])
    barrier_idx = np.arange(nodes, n, dtype=np.int64)
    barrier_val = _barrier_sigma(rng, nodes)
    jac = sp.hstack([_laplacian(k, dim), -sp.identity(nodes, format="csr")], format="csr")
    rhs = rng.standard_normal(n + m)
    return assemble_kkt(f"ocp{dim}d_distributed_k{k}", n, m, hess_idx, hess_idx, hess_val, barrier_idx, barrier_val, jac, rhs)


def ocp_boundary_3d(k: int, alpha: float = 1e-3, seed: int = 0) -> KKTSystem:
    """3-D boundary control (config 5): states on k^3 nodes, one control per boundary-face node
    (6 k^2 controls), -Lap_h y - B u = f with B injecting each face control into its node."""
    rng = np.random.default_rng(seed)
    nodes = k ** 3
    idx = np.arange(nodes).reshape(k, k, k)
    faces = [idx[0, :, :], idx[-1, :, :], idx[:, 0, :], idx[:, -1, :], idx[:, :, 0], idx[:, :, -1]]
    face_nodes = np.concatenate([f.ravel() for f in faces])
    nu = face_nodes.size
    n, m = nodes + nu, nodes
    hess_idx = np.arange(n, dtype=np.int64)
    hess_val = np.concatenate([np.ones(nodes), alpha * np.ones(nu)])
    barrier_idx = np.arange(nodes, n, dtype=np.int64)
    barrier_val = _barrier_sigma(rng, nu)
    inj = sp.coo_matrix((-(k + 1.0) * np.ones(nu), (face_nodes, np.arange(nu))), shape=(nodes, nu)).tocsr()
    jac = sp.hstack([_laplacian(k, 3), inj], format="csr")
    rhs = rng.standard_normal(n + m)
    return assemble_kkt(f"ocp3d_boundary_k{k}", n, m, hess_idx, hess_idx, hess_val, barrier_idx, barrier_val, jac, rhs)


def opf_like(buses: int = 10000, branches: int = 13500, gens: int = 2000, seed: int = 0, duplicate_rows: int = 0,
             delta_w: float = 0.0, curvature: float = 1.0, barrier_exp=(-8.0, 8.0)) -> KKTSystem:
    """AC-OPF-like KKT on a random connected network (config 3).  Variables: (V, theta) per bus,
    (Pg, Qg) per generator, one slack per flow limit (2 per branch).  Constraints: 2 balance rows
    per bus, 2 flow rows per branch.  The Hessian has an indefinite 4x4 block per branch (scaled
    by random multipliers), so the inertia is generally wrong without regularization.
    duplicate_rows > 0 copies that many constraint rows (rank-deficient Jacobian -> zero pivots);
    curvature scales the indefinite branch blocks (large values leave negative curvature in the
    null space of the Jacobian, which the caller's inertia-correction loop has to regularize away)."""
    rng = np.random.default_rng(seed)
    # random spanning tree + extra edges
    parents = np.array([rng.integers(0, i) for i in range(1, buses)], dtype=np.int64)
    tree = np.stack([parents, np.arange(1, buses, dtype=np.int64)], axis=1)
    extra_needed = branches - (buses - 1)
    extra = set()
    while len(extra) < extra_needed:
        a, b = rng.integers(0, buses, size=2)
        if a != b:
            extra.add((int(min(a, b)), int(max(a, b))))
    extra = np.array(sorted(extra), dtype=np.int64).reshape(-1, 2)
    edges = np.concatenate([tree, extra], axis=0)
    nb, ne = buses, edges.shape[0]
    gen_bus = np.sort(rng.choice(buses, size=gens, replace=False))
    # variable layout
    iV = np.arange(nb)
    iT = nb + np.arange(nb)
    iPg = 2 * nb + np.arange(gens)
    iQg = 2 * nb + gens + np.arange(gens)
    iS = 2 * nb + 2 * gens + np.arange(2 * ne)
    n = 2 * nb + 2 * gens + 2 * ne
    # Hessian: per branch a symmetric 4x4 block on (V_a, V_b, T_a, T_b), entries U(-1,1) * y
    hr, hc, hv = [], [], []
    ymul = rng.standard_normal(ne)
    for e in range(ne):
        a, b = edges[e]
        vs = np.sort(np.array([iV[a], iV[b], iT[a], iT[b]]))
        blk = rng.uniform(-1.0, 1.0, size=(4, 4)) * (ymul[e] * curvature)
        for q in range(4):          # column by column, upper triangle
            for p in range(q + 1):
                hr.append(vs[p]); hc.append(vs[q]); hv.append(blk[p, q])
    # generator cost curvature
    for g in range(gens):
        hr.append(iPg[g]); hc.append(iPg[g]); hv.append(rng.uniform(0.01, 0.1))
    hr = np.array(hr, dtype=np.int64); hc = np.array(hc, dtype=np.int64); hv = np.array(hv)
    order = np.lexsort((hr, hc))   # column-major like ASL
    hr, hc, hv = hr[order], hc[order], hv[order]
    # barrier terms on V, Pg, Qg, slacks
    barrier_idx = np.concatenate([iV, iPg, iQg, iS]).astype(np.int64)
    barrier_idx.sort()
    barrier_val = _barrier_sigma(rng, barrier_idx.size, *barrier_exp)
    # Jacobian
    jr, jc, jv = [], [], []
    inc = [[] for _ in range(nb)]
    for e in range(ne):
        inc[edges[e, 0]].append(e); inc[edges[e, 1]].append(e)
    gen_at = {int(b): g for g, b in enumerate(gen_bus)}
    for bus in range(nb):
        for part, igen in ((0, iPg), (1, iQg)):
            r = 2 * bus + part
            nbrs = {bus}
            for e in inc[bus]:
                nbrs.add(int(edges[e, 0])); nbrs.add(int(edges[e, 1]))
            for q in sorted(nbrs):
                jr += [r, r]; jc += [iV[q], iT[q]]; jv += list(rng.uniform(-10, 10, size=2))
            if bus in gen_at:
                jr.append(r); jc.append(igen[gen_at[bus]]); jv.append(-1.0)
    for e in range(ne):
        a, b = edges[e]
        for part in range(2):
            r = 2 * nb + 2 * e + part
            for q in (a, b):
                jr += [r, r]; jc += [iV[q], iT[q]]; jv += list(rng.uniform(-10, 10, size=2))
            jr.append(r); jc.append(iS[2 * e + part]); jv.append(-1.0)
    m = 2 * nb + 2 * ne
    jac = sp.coo_matrix((jv, (jr, jc)), shape=(m, n)).tocsr()
    if duplicate_rows > 0:
        dup = jac[:duplicate_rows, :]
        jac = sp.vstack([jac, dup], format="csr")
        m += duplicate_rows
    rhs = rng.standard_normal(n + m)
    kkt = assemble_kkt(f"opf_like_b{buses}", n, m, hr, hc, hv, barrier_idx, barrier_val, jac, rhs)
    if delta_w != 0.0:
        kkt.val = kkt.with_regularization(delta_w, 0.0)
    return kkt


def kat_matrix():
    """5x5 known-answer matrix of the reference's dormant solver tests
    (unotest/functional_tests/MUMPSSolverTests.cpp:11-61): x = (1..5), inertia (3,2,0)."""
    row = np.array([0, 0, 1, 1, 2, 2, 4], dtype=np.uint64)
    col = np.array([0, 1, 2, 4, 2, 3, 4], dtype=np.uint64)
    val = np.array([2.0, 3.0, 4.0, 6.0, 1.0, 5.0, 1.0])
    rhs = np.array([8.0, 45.0, 31.0, 15.0, 17.0])
    return KKTSystem("kat5", 5, 0, 0, row, col, val, rhs)


def kat_singular():
    """4x4 singular matrix with duplicate entries (MUMPSSolverTests.cpp:63-83): inertia (1,1,2)."""
    row = np.array([0, 0, 0, 1, 1, 2, 3], dtype=np.uint64)
    col = np.array([0, 0, 1, 1, 1, 2, 3], dtype=np.uint64)
    val = np.array([-0.0198, 0.625075, -0.277512, -0.624975, 0.625075, 0.0, 0.0])
    return KKTSystem("kat_singular4", 4, 0, 0, row, col, val, np.ones(4))
