"""SURVEY 8(f) ranks 1 and 3 on the GPU against restatements of the reference's host loops.

rank 1: the KKT value array generated on the device (slots, Hessian, barrier diagonals, Jacobian) and the device-side diagonal
shift must give BIT-IDENTICAL factors, inertia and solution to the host-assembled path (b200kkt_factorize / refactorize_shifted).
rank 3: right-hand side, K x and the direction unpack with both fraction-to-boundary rules, bit for bit against numpy code that
performs the reference's operations in the reference's order (np.add.at accumulates sequentially, like the host loops)."""
import numpy as np
import pytest

from uno_b200 import KKTSolver
from uno_b200 import generators as G

pytestmark = pytest.mark.gpu


def ipm_state(kkt, seed=0):
    """A synthetic interior-point iterate for an ocp_distributed system: states free, controls in [-1, 1]."""
    rng = np.random.default_rng(seed)
    n, m = kkt.n, kkt.m
    nodes = n // 2
    lower = np.full(n, -np.inf)
    upper = np.full(n, np.inf)
    lower[nodes:] = -1.0
    upper[nodes:] = 1.0
    x = rng.standard_normal(n)
    x[nodes:] = rng.uniform(-0.999, 0.999, size=nodes)
    zl = np.where(np.isfinite(lower), 10.0 ** rng.uniform(-6, 2, size=n), 0.0)
    zu = np.where(np.isfinite(upper), -(10.0 ** rng.uniform(-6, 2, size=n)), 0.0)    # upper-bound duals are non-positive in Uno
    layout = (n, m, n, nodes, kkt.nnz - (n + m) - n - nodes)      # slots | Hessian diag (n entries) | barrier (controls) | Jacobian
    barrier_var = np.arange(nodes, n, dtype=np.int32)
    return dict(lower=lower, upper=upper, x=x, zl=zl, zu=zu, layout=layout, barrier_var=barrier_var, rng=rng)


def host_values(kkt, st, delta_w, delta_c):
    """Subproblem.cpp:57-70 + PrimalDualInteriorPointProblem.cpp:56-78 + COOFormat.hpp:101-110 on the host."""
    n, m = kkt.n, kkt.m
    npri, ndual, nh, nb, nj = st["layout"]
    v = kkt.val.copy()
    v[:npri] = delta_w
    v[npri:npri + ndual] = -delta_c
    i = st["barrier_var"]
    t = np.zeros(nb)
    t = t + st["zl"][i] / (st["x"][i] - st["lower"][i])      # both bounds finite for the controls
    t = t + st["zu"][i] / (st["x"][i] - st["upper"][i])
    o = npri + ndual + nh
    v[o:o + nb] = t
    return v


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_device_generated_kkt_values_factorize_bit_identically():
    import torch
    kkt = G.ocp_distributed(20, 2, seed=3)
    st = ipm_state(kkt)
    npri, ndual, nh, nb, nj = st["layout"]
    o = npri + ndual
    d = {k: dev(torch, st[k]) for k in ("x", "lower", "upper", "zl", "zu", "barrier_var")}
    d_h = dev(torch, kkt.val[o:o + nh])
    d_j = dev(torch, kkt.val[o + nh + nb:])
    d_rhs = dev(torch, kkt.rhs)
    for dw, dc in [(0.0, 0.0), (1e-4, 1e-8)]:
        ref = KKTSolver(device=0)
        ref.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
        vals = host_values(kkt, st, dw, dc)
        ref.do_numerical_factorization(vals)
        s = KKTSolver(device=0)
        s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
        s.assemble_kkt_values_device(st["layout"], d_h.data_ptr(), d["barrier_var"].data_ptr(), d["x"].data_ptr(), d["lower"].data_ptr(), d["upper"].data_ptr(),
                                     d["zl"].data_ptr(), d["zu"].data_ptr(), d_j.data_ptr(), dw, dc)
        s.factorize_resident()
        assert s.get_inertia() == ref.get_inertia()
        assert all(np.array_equal(a, b) for a, b in zip(ref.debug_factor(), s.debug_factor()))
        d_x = torch.empty_like(d_rhs)
        s.solve_device(d_rhs.data_ptr(), d_x.data_ptr())
        assert np.array_equal(d_x.cpu().numpy(), ref.solve_indefinite_system(kkt.rhs))
        # the inertia-correction step with no PCIe traffic against the host fast path
        ref.refactorize_shifted(host_values(kkt, st, 100.0 * max(dw, 1e-4), 1e-8)[: kkt.reg_size])
        s.shift_diagonal_device(npri, 100.0 * max(dw, 1e-4), 1e-8)
        s.factorize_resident()
        assert s.get_inertia() == ref.get_inertia()
        assert all(np.array_equal(a, b) for a, b in zip(ref.debug_factor(), s.debug_factor()))


def test_rhs_product_and_direction_match_the_reference_loops_bit_for_bit():
    import torch
    kkt = G.ocp_distributed(24, 2, seed=5)
    st = ipm_state(kkt, seed=1)
    rng = st["rng"]
    n, m = kkt.n, kkt.m
    s = KKTSolver(device=0)
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
    vals = host_values(kkt, st, 0.0, 0.0)
    s.do_numerical_factorization(vals)          # makes the values resident
    row, col = kkt.row.astype(np.int64), kkt.col.astype(np.int64)

    # ---- Subproblem::assemble_augmented_rhs (Subproblem.cpp:80-99) ------------------------------------------------------------
    g = rng.standard_normal(n)
    c = rng.standard_normal(m)
    y = rng.standard_normal(m)
    y[::7] = 0.0                                # the reference skips zero multipliers
    rhs = np.zeros(n + m)
    rhs[:n] = -g
    jac = np.flatnonzero((row < n) & (col >= n))                      # ascending slot = constraint by constraint, entry by entry
    jy = y[col[jac] - n]
    keep = jy != 0.0
    np.add.at(rhs, row[jac][keep], jy[keep] * vals[jac][keep])
    rhs[n:] = -c
    d_rhs = torch.empty(n + m, dtype=torch.float64, device="cuda")
    d_g, d_c, d_y = dev(torch, g), dev(torch, c), dev(torch, y)          # keep the tensors alive: their memory is what the call reads
    s.assemble_rhs_device(d_g.data_ptr(), d_c.data_ptr(), d_y.data_ptr(), d_rhs.data_ptr())
    assert np.array_equal(d_rhs.cpu().numpy(), rhs)

    # ---- SymmetricMatrix::product (SymmetricMatrix.hpp:99-109) ------------------------------------------------------------------
    xv = rng.standard_normal(n + m)
    off = row != col
    tgt = np.empty(2 * kkt.nnz, dtype=np.int64)
    contrib = np.empty(2 * kkt.nnz)
    tgt[0::2], contrib[0::2] = row, vals * xv[col]
    tgt[1::2], contrib[1::2] = col, vals * xv[row]
    mask = np.ones(2 * kkt.nnz, dtype=bool)
    mask[1::2] = off
    prod = np.zeros(n + m)
    np.add.at(prod, tgt[mask], contrib[mask])
    d_prod = torch.empty(n + m, dtype=torch.float64, device="cuda")
    d_xv = dev(torch, xv)
    s.symmetric_product_device(d_xv.data_ptr(), d_prod.data_ptr())
    assert np.array_equal(d_prod.cpu().numpy(), prod)

    # ---- assemble_primal_dual_direction (PrimalDualInteriorPointProblem.cpp:173-194, 262-325) --------------------------------
    mu, tau = 0.1, max(0.99, 1.0 - 0.1)
    sol = s.solve_indefinite_system(rhs)
    x, lo, up, zl, zu = st["x"], st["lower"], st["upper"], st["zl"], st["zu"]
    dx = sol[:n].copy()
    dy = -sol[n:]
    fl, fu = np.isfinite(lo), np.isfinite(up)
    with np.errstate(invalid="ignore", divide="ignore"):
        dzl = np.where(fl, (mu - dx * zl) / (x - lo) - zl, 0.0)
        dzu = np.where(fu, (mu - dx * zu) / (x - up) - zu, 0.0)
        cand_p = np.concatenate([(-tau * (x - lo) / dx)[fl & (dx < 0)], (-tau * (x - up) / dx)[fu & (dx > 0)]])
        cand_d = np.concatenate([(-tau * zl / dzl)[fl & (dzl < 0)], (-tau * zu / dzu)[fu & (dzu > 0)]])
    ps = min([1.0] + [t for t in cand_p if t > 0])
    ds = min([1.0] + [t for t in cand_d if t > 0])
    out = {k: torch.empty(n if k != "dy" else m, dtype=torch.float64, device="cuda") for k in ("dx", "dy", "dzl", "dzu")}
    d_in = {k: dev(torch, v) for k, v in dict(sol=sol, x=x, lo=lo, up=up, zl=zl, zu=zu).items()}
    gps, gds = s.unpack_direction_device(mu, tau, d_in["sol"].data_ptr(), d_in["x"].data_ptr(), d_in["lo"].data_ptr(), d_in["up"].data_ptr(),
                                         d_in["zl"].data_ptr(), d_in["zu"].data_ptr(), out["dx"].data_ptr(), out["dy"].data_ptr(), out["dzl"].data_ptr(),
                                         out["dzu"].data_ptr())
    assert (gps, gds) == (ps, ds) and 0.0 < ps <= 1.0 and 0.0 < ds <= 1.0
    assert np.array_equal(out["dx"].cpu().numpy(), dx * ps) and np.array_equal(out["dy"].cpu().numpy(), dy * ps)
    assert np.array_equal(out["dzl"].cpu().numpy(), dzl * ds) and np.array_equal(out["dzu"].cpu().numpy(), dzu * ds)
