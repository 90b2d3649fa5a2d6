"""Host-side logic through the C ABI without a GPU: library symbols, symbolic phase, plan
invariants, the assembly permutation, error behaviour.  No numeric call can succeed here."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

from emulator import EmulatedFactorization
from helpers import backward_error, kat_system
from uno_b200 import B200KKTError, KKTSolver, capi
from uno_b200 import generators as G

ROOT = Path(__file__).resolve().parent.parent


def host_solver(kkt, **opts):
    s = KKTSolver(device=-1, **opts)
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n if kkt.m else 0, values=kkt.val)
    return s


def test_library_exports_every_declared_symbol(lib):
    header = (ROOT / "include" / "b200kkt.h").read_text()
    declared = set(re.findall(r"\b(b200kkt_[a-z_]+)\s*\(", header))
    assert declared == set(capi.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name


def test_numeric_calls_fail_loudly_without_device():
    kkt = G.kat_matrix()
    s = host_solver(kkt)
    with pytest.raises(B200KKTError, match="no CUDA device"):
        s.do_numerical_factorization(kkt.val)
    with pytest.raises(B200KKTError):
        s.solve_indefinite_system(kkt.rhs)
    with pytest.raises(B200KKTError):
        s.get_inertia()


def test_bad_input_is_rejected():
    s = KKTSolver(device=-1)
    with pytest.raises(B200KKTError, match="below the diagonal"):
        s.do_symbolic_analysis(3, [2], [0])
    with pytest.raises(B200KKTError, match="out of range"):
        s.do_symbolic_analysis(3, [0], [3])
    with pytest.raises(B200KKTError, match="unknown option"):
        s.set_option("no_such_option", 1)
    with pytest.raises(B200KKTError):
        s.permutation()


def check_plan_invariants(kkt, s):
    plan = s.plan()
    perm, scat = s.permutation()
    n = kkt.dim
    assert sorted(perm.tolist()) == list(range(n))
    k, f, ld, off, parent, first = plan["k"], plan["f"], plan["ld"], plan["pool_offset"], plan["parent"], plan["first_col"]
    nsn = plan["num_fronts"]
    assert k.sum() == n and np.all(k >= 1) and np.all(f >= k) and np.all(ld >= f) and np.all(ld % 4 == 0)
    assert np.all(first == np.concatenate([[0], np.cumsum(k)[:-1]]))
    assert np.all(off[1:] == off[:-1] + ld[:-1].astype(np.int64) * k[:-1]) and off[-1] + ld[-1] * k[-1] == plan["pool_doubles"]
    rows_off, rows = plan["rows_off"], plan["rows"]
    for sn in range(nsn):
        r = rows[rows_off[sn] : rows_off[sn + 1]]
        assert len(r) == f[sn]
        assert np.all(r[: k[sn]] == first[sn] + np.arange(k[sn]))
        b = r[k[sn] :]
        assert np.all(np.diff(b) > 0) and (len(b) == 0 or b[0] >= first[sn] + k[sn])
        if len(b):
            p = parent[sn]
            assert p > sn and first[p] <= b[0] < first[p] + k[p]
            pr = set(rows[rows_off[p] : rows_off[p + 1]].tolist())
            assert set(b.tolist()) <= pr          # child boundary is contained in the parent front
        else:
            assert parent[sn] == -1
    # assembly permutation: recompute every destination independently from the front row lists
    iperm = np.empty(n, dtype=np.int64)
    iperm[perm] = np.arange(n)
    sn_of = np.repeat(np.arange(nsn), k)
    pr, pc = iperm[kkt.row.astype(np.int64)], iperm[kkt.col.astype(np.int64)]
    lo, hi = np.minimum(pr, pc), np.maximum(pr, pc)
    for e in range(kkt.nnz):
        sn = sn_of[lo[e]]
        r = rows[rows_off[sn] : rows_off[sn + 1]]
        lrow = int(np.searchsorted(r[k[sn] :], hi[e])) + k[sn] if hi[e] >= first[sn] + k[sn] else hi[e] - first[sn]
        assert r[lrow] == hi[e]
        assert scat[e] == off[sn] + (lo[e] - first[sn]) * ld[sn] + lrow
    return plan, perm, scat


@pytest.mark.parametrize("make", [G.kat_matrix, G.kat_singular, lambda: G.ocp_distributed(9, 2), lambda: G.ocp_distributed(5, 3),
                                  lambda: G.ocp_boundary_3d(5), lambda: G.opf_like(30, 40, 6, seed=3)])
def test_plan_invariants_and_assembly_permutation(make):
    kkt = make()
    check_plan_invariants(kkt, host_solver(kkt))


@pytest.mark.parametrize("opts", [{}, {"ordering": 1}, {"kkt_pairing": 0}, {"relax_small": 1, "relax_mid": 1, "relax_big": 1, "relax_any_zeros": 0.0}])
def test_plan_drives_a_correct_factorization(opts):
    """The numpy emulation of the device algorithm, run on the plan, solves the system and finds the inertia."""
    kkt = G.ocp_distributed(10, 2, seed=2)
    s = host_solver(kkt, **opts)
    plan, perm, scat = check_plan_invariants(kkt, s)
    emu = EmulatedFactorization(plan, perm, scat, kkt, kkt.val)
    assert emu.inertia() == (kkt.n, kkt.m, 0)
    assert backward_error(kkt, kkt.val, emu.solve(kkt.rhs)) < 1e-12


def test_reference_kats_through_plan_emulation():
    import json
    kats = json.loads((ROOT / "tests/golden/reference_kats.json").read_text())
    e = kats["kat1_solve"]
    kkt = kat_system(e)
    s = host_solver(kkt)
    perm, scat = s.permutation()
    emu = EmulatedFactorization(s.plan(), perm, scat, kkt, kkt.val)
    assert np.max(np.abs(emu.solve(kkt.rhs) - np.array(e["x"]))) < e["tol"]
    assert list(emu.inertia()) == kats["kat2_inertia"]["inertia"]
    kkt3 = kat_system(kats["kat3_singular"])
    s3 = host_solver(kkt3)
    perm, scat = s3.permutation()
    assert list(EmulatedFactorization(s3.plan(), perm, scat, kkt3, kkt3.val).inertia()) == kats["kat3_singular"]["inertia"]


def test_user_permutation_and_reanalysis_with_another_pattern():
    # Preprocessing.cpp:60 analyses [[I, J^T],[J, 0]] (no regularization slots) on the same solver
    # instance before PrimalDualRegularization.hpp:147 analyses the KKT pattern.
    kkt = G.ocp_distributed(6, 2)
    s = KKTSolver(device=-1)
    rng = np.random.default_rng(1)
    s.set_user_permutation(rng.permutation(kkt.dim))
    s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n)
    check_plan_invariants(kkt, s)
    s.set_user_permutation(None)
    sub = slice(kkt.reg_size, None)
    kkt2 = type(kkt)("noreg", kkt.n, kkt.m, 0, kkt.row[sub], kkt.col[sub], kkt.val[sub], kkt.rhs)
    s.do_symbolic_analysis(kkt2.dim, kkt2.row, kkt2.col, 0, kkt2.n)
    check_plan_invariants(kkt2, s)
    assert s.stats()["reg_size"] == 0


def test_stats_work_figures_are_consistent():
    kkt = G.ocp_distributed(12, 2)
    s = host_solver(kkt)
    st, plan = s.stats(), s.plan()
    k, f = plan["k"].astype(np.float64), plan["f"].astype(np.float64)
    r = f - k
    assert st["nnz_factor"] == int(np.sum(k * f - k * (k - 1) / 2))
    assert np.isclose(st["flops_factor"], np.sum(k ** 3 / 3 + k * k * r + k * r * r))
    assert st["num_pairs"] == kkt.m and st["assembly_distinct"] == plan["assembly_distinct"] <= kkt.nnz


def test_stats_struct_layout_matches_header():
    header = (ROOT / "include" / "b200kkt.h").read_text()
    body = header[header.index("typedef struct b200kkt_stats {") : header.index("} b200kkt_stats;")]
    names = re.findall(r"\b([a-z_0-9]+)\s*[,;]", re.sub(r"/\*.*?\*/", "", body, flags=re.S))
    assert names == [n for n, _ in capi.Stats._fields_]
    assert ctypes.sizeof(capi.Stats) == 8 * len(names)


def test_process_parallel_ordering_is_a_deterministic_permutation_of_serial_quality(lib):
    """Big graphs: the top levels of the nested dissection are split by the library and their halves ordered by forked processes
    (METIS has process-wide state, so threads would make the ordering depend on the interleaving).  The result must be a
    permutation, the same on every run - every rank of a partitioned run computes it on its own - and as good as METIS_NodeND's."""
    from uno_b200 import KKTSolver
    from uno_b200 import generators as G
    kkt = G.ocp_distributed(40, 3, seed=0)                       # 128 000 nodes after KKT pairing
    runs = []
    for depth in (3, 3, 0):
        s = KKTSolver(device=-1, parallel_ordering_depth=depth, parallel_ordering_min_nodes=50000)
        s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
        perm, _ = s.permutation()
        st = s.stats()
        runs.append((perm.copy(), st["flops_factor"], s.plan()["ordering"] if "ordering" in s.plan() else None))
        s.close()
    assert np.array_equal(np.sort(runs[0][0]), np.arange(kkt.dim))
    assert np.array_equal(runs[0][0], runs[1][0])
    assert not np.array_equal(runs[0][0], runs[2][0])            # the serial path is another ordering ...
    assert runs[0][1] <= 1.15 * runs[2][1]                       # ... of the same quality
