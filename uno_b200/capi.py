"""ctypes binding of include/b200kkt.h plus a thin test/bench-side mirror of the reference's
solver interface (uno/ingredients/subproblem_solvers/DirectSymmetricIndefiniteLinearSolver.hpp:13-25).

The product is the shared library; this file only marshals numpy arrays into it.  There is no
fallback: if the library is missing or the handle has no CUDA device, numeric calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_LIB_PATH = Path(os.environ["B200KKT_LIB"]) if os.environ.get("B200KKT_LIB") else Path(__file__).resolve().parent / "lib" / "libb200kkt.so"   # B200KKT_LIB: instrumented build variants (profiling)

EXPORTED_SYMBOLS = [
    "b200kkt_create", "b200kkt_destroy", "b200kkt_set_option", "b200kkt_set_user_permutation", "b200kkt_analyse",
    "b200kkt_factorize", "b200kkt_refactorize_shifted", "b200kkt_factorize_device", "b200kkt_inertia", "b200kkt_solve",
    "b200kkt_solve_device", "b200kkt_get_permutation", "b200kkt_get_stats", "b200kkt_last_error", "b200kkt_plan_sizes",
    "b200kkt_plan_fronts", "b200kkt_plan_rows", "b200kkt_debug_assemble", "b200kkt_debug_factor", "b200kkt_get_stream",
    "b200kkt_debug_bench_update", "b200kkt_debug_phase_cycles", "b200kkt_debug_bench_bk",
    "b200kkt_comm_unique_id", "b200kkt_comm_init", "b200kkt_plan_partition", "b200kkt_debug_delay_columns",
    "b200kkt_backward_error_noalloc", "b200kkt_set_pairing_values",
    "b200kkt_assemble_kkt_values_device", "b200kkt_shift_diagonal_device", "b200kkt_factorize_resident", "b200kkt_get_values_device",
    "b200kkt_assemble_rhs_device", "b200kkt_symmetric_product_device", "b200kkt_unpack_direction_device",
]


class KKTLayout(C.Structure):
    _fields_ = [("num_primal_slots", C.c_uint64), ("num_dual_slots", C.c_uint64), ("nnz_hessian", C.c_uint64), ("num_barrier", C.c_uint64),
                ("nnz_jacobian", C.c_uint64)]


class Stats(C.Structure):
    _fields_ = [
        ("n", C.c_int64), ("nnz", C.c_int64), ("reg_size", C.c_int64),
        ("num_fronts", C.c_int64), ("num_levels", C.c_int64), ("max_front", C.c_int64), ("max_pivots", C.c_int64),
        ("num_pairs", C.c_int64), ("nnz_factor", C.c_int64), ("pool_doubles", C.c_int64), ("assembly_distinct", C.c_int64),
        ("flops_factor", C.c_double), ("flops_tensor", C.c_double),
        ("t_ordering", C.c_double), ("t_symbolic", C.c_double), ("t_maps", C.c_double), ("t_upload", C.c_double),
        ("t_factor_last", C.c_double), ("t_solve_last", C.c_double),
        ("refine_steps_last", C.c_int64), ("backward_error_last", C.c_double), ("kernel_launches", C.c_int64),
        ("num_null_pivots_last", C.c_int64), ("num_two_by_two_last", C.c_int64),
        ("t_update_profiled", C.c_double), ("flops_update_profiled", C.c_double), ("n_update_profiled", C.c_int64),
        ("t_assemble_profiled", C.c_double),
        ("num_parts", C.c_int64), ("part_rank", C.c_int64), ("num_cut_fronts", C.c_int64),
        ("flops_own", C.c_double), ("flops_critical", C.c_double),
        ("comm_bytes_factor", C.c_int64), ("comm_bytes_sweep", C.c_int64),
        ("t_shared_profiled", C.c_double), ("flops_shared_profiled", C.c_double), ("t_shared_stall_profiled", C.c_double),
        ("t_before_shared_profiled", C.c_double), ("t_transfer_wait_profiled", C.c_double),
        ("num_delayed_total", C.c_int64), ("num_replans_total", C.c_int64), ("num_replans_last", C.c_int64),
        ("backward_error_componentwise_last", C.c_double),
        ("cb_arena_doubles", C.c_int64), ("pool_doubles_global", C.c_int64), ("cb_arena_doubles_global", C.c_int64),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


_lib = None


def load_library():
    """Load libb200kkt.so (built in-tree by uno_b200.build).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise RuntimeError(f"{_LIB_PATH} is missing: run `python -m uno_b200.build` (there is no fallback path)")
    lib = C.CDLL(str(_LIB_PATH))
    H = C.c_void_p
    P = C.POINTER
    lib.b200kkt_create.argtypes = [P(H), C.c_int, C.c_int]
    lib.b200kkt_destroy.argtypes = [H]
    lib.b200kkt_set_option.argtypes = [H, C.c_char_p, C.c_double]
    lib.b200kkt_set_user_permutation.argtypes = [H, C.c_void_p, C.c_uint64]
    lib.b200kkt_analyse.argtypes = [H, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.b200kkt_factorize.argtypes = [H, C.c_void_p]
    lib.b200kkt_refactorize_shifted.argtypes = [H, C.c_void_p]
    lib.b200kkt_factorize_device.argtypes = [H, C.c_void_p]
    lib.b200kkt_inertia.argtypes = [H, P(C.c_uint64), P(C.c_uint64), P(C.c_uint64)]
    lib.b200kkt_solve.argtypes = [H, C.c_void_p, C.c_void_p, P(C.c_double)]
    lib.b200kkt_solve_device.argtypes = [H, C.c_void_p, C.c_void_p, P(C.c_double)]
    lib.b200kkt_get_permutation.argtypes = [H, C.c_void_p, C.c_void_p]
    lib.b200kkt_get_stats.argtypes = [H, P(Stats)]
    lib.b200kkt_last_error.argtypes = [H]
    lib.b200kkt_last_error.restype = C.c_char_p
    lib.b200kkt_get_stream.argtypes = [H, P(C.c_void_p)]
    lib.b200kkt_plan_sizes.argtypes = [H, P(C.c_int64)]
    lib.b200kkt_plan_fronts.argtypes = [H] + [C.c_void_p] * 7
    lib.b200kkt_plan_rows.argtypes = [H, C.c_void_p, C.c_void_p]
    lib.b200kkt_debug_assemble.argtypes = [H, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.b200kkt_debug_factor.argtypes = [H] + [C.c_void_p] * 5
    lib.b200kkt_debug_delay_columns.argtypes = [H, C.c_uint64, C.c_void_p]
    lib.b200kkt_set_pairing_values.argtypes = [H, C.c_void_p, C.c_uint64]
    VP = C.c_void_p
    lib.b200kkt_assemble_kkt_values_device.argtypes = [H, P(KKTLayout), VP, VP, VP, VP, VP, VP, VP, VP, C.c_double, C.c_double]
    lib.b200kkt_shift_diagonal_device.argtypes = [H, C.c_uint64, C.c_double, C.c_double]
    lib.b200kkt_factorize_resident.argtypes = [H]
    lib.b200kkt_get_values_device.argtypes = [H, P(C.c_void_p)]
    lib.b200kkt_assemble_rhs_device.argtypes = [H, VP, VP, VP, VP]
    lib.b200kkt_symmetric_product_device.argtypes = [H, VP, VP]
    lib.b200kkt_unpack_direction_device.argtypes = [H, C.c_double, C.c_double] + [VP] * 10 + [P(C.c_double), P(C.c_double)]
    lib.b200kkt_backward_error_noalloc.argtypes = [H, C.c_void_p, C.c_void_p, P(C.c_double), P(C.c_double)]
    lib.b200kkt_comm_unique_id.argtypes = [C.c_void_p]
    lib.b200kkt_comm_init.argtypes = [H, C.c_int, C.c_int, C.c_void_p]
    lib.b200kkt_plan_partition.argtypes = [H, C.c_void_p, C.c_void_p, P(C.c_int64), C.c_void_p, C.c_void_p, P(C.c_double)]
    for name in EXPORTED_SYMBOLS:
        if name != "b200kkt_last_error":
            getattr(lib, name).restype = C.c_int
    _lib = lib
    return lib


class B200KKTError(RuntimeError):
    pass


def comm_unique_id() -> bytes:
    """128-byte NCCL unique id (rank 0 creates it, the host program hands it to the other ranks)."""
    lib = load_library()
    buf = C.create_string_buffer(128)
    rc = lib.b200kkt_comm_unique_id(buf)
    if rc != 0:
        raise B200KKTError(f"b200kkt_comm_unique_id failed ({rc}): libnccl.so.2 could not be opened")
    return buf.raw


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class KKTSolver:
    """Python mirror of the reference solver interface, calling the C ABI.

    Method names follow DirectSymmetricIndefiniteLinearSolver / MUMPSSolver
    (uno/ingredients/subproblem_solvers/MUMPS/MUMPSSolver.hpp:17-40).  The matrix is passed as the
    reference's COO triple (row <= col, 0-based, duplicates summed, leading regularization slots).
    """

    def __init__(self, device: int = 0, num_gpus: int = 1, **options):
        self._lib = load_library()
        self._h = C.c_void_p()
        rc = self._lib.b200kkt_create(C.byref(self._h), device, num_gpus)
        if rc != 0:
            msg = self._lib.b200kkt_last_error(self._h).decode() if self._h else "create failed"
            if self._h:
                self._lib.b200kkt_destroy(self._h)
                self._h = C.c_void_p()
            raise B200KKTError(f"b200kkt_create failed ({rc}): {msg}")
        self.n = 0
        self.nnz = 0
        self.reg_size = 0
        for k, v in options.items():
            self.set_option(k, v)

    # -- life cycle --------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.b200kkt_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise B200KKTError(f"{what} failed ({rc}): {self._lib.b200kkt_last_error(self._h).decode()}")

    def set_option(self, name: str, value: float):
        self._check(self._lib.b200kkt_set_option(self._h, name.encode(), float(value)), "set_option")

    def set_user_permutation(self, perm):
        if perm is None:
            self._check(self._lib.b200kkt_set_user_permutation(self._h, None, 0), "set_user_permutation")
            return
        p = np.ascontiguousarray(perm, dtype=np.int64)
        self._check(self._lib.b200kkt_set_user_permutation(self._h, _ptr(p), p.size), "set_user_permutation")

    # -- multi-GPU (one process per GPU) ---------------------------------------------------------
    def comm_init(self, rank: int, world: int, unique_id: bytes):
        """Join the NCCL communicator of a subtree-partitioned solve; `unique_id` = rank 0's comm_unique_id()."""
        assert len(unique_id) == 128
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._check(self._lib.b200kkt_comm_init(self._h, int(rank), int(world), buf), "comm_init")

    def partition(self):
        """Subtree map of the last analysis: owner per front, cut fronts, factor flops per rank, critical-path flops."""
        nsn = self.plan()["num_fronts"]
        owner = np.empty(nsn, dtype=np.int32)
        shared = np.empty(nsn, dtype=np.int8)
        cut = np.empty(nsn, dtype=np.int32)
        ncut = C.c_int64()
        parts = np.zeros(max(1, int(self.stats()["num_parts"])), dtype=np.float64)
        crit = C.c_double()
        self._check(self._lib.b200kkt_plan_partition(self._h, _ptr(owner), _ptr(shared), C.byref(ncut), _ptr(cut), _ptr(parts), C.byref(crit)), "plan_partition")
        return {"owner": owner, "shared": shared.astype(bool), "cut_fronts": cut[:ncut.value].copy(), "part_flops": parts, "critical_flops": crit.value}

    # -- reference interface -----------------------------------------------------------------
    def initialize_memory(self, number_variables, number_constraints, number_hessian_nonzeros, regularization_size):
        """MUMPSSolver::initialize_memory (MUMPSSolver.cpp:51-70): sizes only; device memory is
        allocated at analysis time, when the front sizes are known."""
        self.number_variables = int(number_variables)
        self.number_constraints = int(number_constraints)
        self.dimension = self.number_variables + self.number_constraints
        self.capacity = int(number_hessian_nonzeros) + int(regularization_size)
        self.reg_size = int(regularization_size)

    def do_symbolic_analysis(self, n, row, col, reg_size=0, num_primal=0, values=None):
        """values: the matrix values at analysis time (the reference's do_symbolic_analysis receives the filled matrix):
        a hint for the KKT pairing only."""
        row = np.ascontiguousarray(row, dtype=np.uint64)
        col = np.ascontiguousarray(col, dtype=np.uint64)
        assert row.shape == col.shape
        if values is not None:
            v = np.ascontiguousarray(values, dtype=np.float64)
            assert v.size == row.size
            self._check(self._lib.b200kkt_set_pairing_values(self._h, _ptr(v), v.size), "set_pairing_values")
        self.n, self.nnz, self.reg_size = int(n), int(row.size), int(reg_size)
        self._check(self._lib.b200kkt_analyse(self._h, self.n, self.nnz, self.reg_size, int(num_primal), _ptr(row), _ptr(col)), "analyse")

    def do_numerical_factorization(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.size == self.nnz
        self._check(self._lib.b200kkt_factorize(self._h, _ptr(v)), "factorize")

    def refactorize_shifted(self, reg_values):
        v = np.ascontiguousarray(reg_values, dtype=np.float64)
        assert v.size == self.reg_size
        self._check(self._lib.b200kkt_refactorize_shifted(self._h, _ptr(v)), "refactorize_shifted")

    def factorize_device(self, device_ptr: int):
        self._check(self._lib.b200kkt_factorize_device(self._h, C.c_void_p(device_ptr)), "factorize_device")

    def get_inertia(self):
        p, n, z = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(self._lib.b200kkt_inertia(self._h, C.byref(p), C.byref(n), C.byref(z)), "inertia")
        return int(p.value), int(n.value), int(z.value)

    def number_negative_eigenvalues(self):
        return self.get_inertia()[1]

    def number_zero_eigenvalues(self):
        return self.get_inertia()[2]

    def matrix_is_singular(self):
        return self.get_inertia()[2] > 0

    def rank(self):
        return self.n - self.get_inertia()[2]

    def solve_indefinite_system(self, rhs, return_backward_error=False):
        b = np.ascontiguousarray(rhs, dtype=np.float64)
        assert b.size == self.n
        x = np.empty(self.n, dtype=np.float64)
        berr = C.c_double()
        self._check(self._lib.b200kkt_solve(self._h, _ptr(b), _ptr(x), C.byref(berr)), "solve")
        return (x, berr.value) if return_backward_error else x

    def solve_device(self, rhs_ptr: int, x_ptr: int):
        berr = C.c_double()
        self._check(self._lib.b200kkt_solve_device(self._h, C.c_void_p(rhs_ptr), C.c_void_p(x_ptr), C.byref(berr)), "solve_device")
        return berr.value

    def backward_error_noalloc(self, x_ptr: int, rhs_ptr: int):
        """(normwise, componentwise) backward error of the device vector x for the last factorized matrix; allocates nothing."""
        nw, cw = C.c_double(), C.c_double()
        self._check(self._lib.b200kkt_backward_error_noalloc(self._h, C.c_void_p(x_ptr), C.c_void_p(rhs_ptr), C.byref(nw), C.byref(cw)),
                    "backward_error_noalloc")
        return nw.value, cw.value

    # -- device-side host steps (all arguments are device pointers as integers) ---------------------------------
    def assemble_kkt_values_device(self, layout, hess_ptr, barrier_var_ptr, x_ptr, lower_ptr, upper_ptr, zl_ptr, zu_ptr, jac_ptr, delta_w=0.0, delta_c=0.0):
        lay = KKTLayout(*[int(v) for v in layout])
        vp = lambda p: C.c_void_p(int(p) if p else None)  # noqa: E731
        self._check(self._lib.b200kkt_assemble_kkt_values_device(self._h, C.byref(lay), vp(hess_ptr), vp(barrier_var_ptr), vp(x_ptr), vp(lower_ptr), vp(upper_ptr),
                                                                 vp(zl_ptr), vp(zu_ptr), vp(jac_ptr), float(delta_w), float(delta_c)), "assemble_kkt_values_device")

    def shift_diagonal_device(self, num_primal_slots, delta_w, delta_c):
        self._check(self._lib.b200kkt_shift_diagonal_device(self._h, int(num_primal_slots), float(delta_w), float(delta_c)), "shift_diagonal_device")

    def factorize_resident(self):
        self._check(self._lib.b200kkt_factorize_resident(self._h), "factorize_resident")

    def values_device(self) -> int:
        p = C.c_void_p()
        self._check(self._lib.b200kkt_get_values_device(self._h, C.byref(p)), "get_values_device")
        return int(p.value or 0)

    def assemble_rhs_device(self, gradient_ptr, constraints_ptr, multipliers_ptr, rhs_ptr):
        self._check(self._lib.b200kkt_assemble_rhs_device(self._h, C.c_void_p(gradient_ptr), C.c_void_p(constraints_ptr), C.c_void_p(multipliers_ptr), C.c_void_p(rhs_ptr)),
                    "assemble_rhs_device")

    def symmetric_product_device(self, x_ptr, y_ptr):
        self._check(self._lib.b200kkt_symmetric_product_device(self._h, C.c_void_p(x_ptr), C.c_void_p(y_ptr)), "symmetric_product_device")

    def unpack_direction_device(self, mu, tau, sol, x, lower, upper, zl, zu, dx, dy, dzl, dzu):
        ps, ds = C.c_double(), C.c_double()
        args = [C.c_void_p(int(p) if p else None) for p in (sol, x, lower, upper, zl, zu, dx, dy, dzl, dzu)]
        self._check(self._lib.b200kkt_unpack_direction_device(self._h, float(mu), float(tau), *args, C.byref(ps), C.byref(ds)), "unpack_direction_device")
        return ps.value, ds.value

    def stream(self) -> int:
        p = C.c_void_p()
        self._check(self._lib.b200kkt_get_stream(self._h, C.byref(p)), "get_stream")
        return int(p.value or 0)

    # -- parity hooks ------------------------------------------------------------------------
    def stats(self):
        s = Stats()
        self._check(self._lib.b200kkt_get_stats(self._h, C.byref(s)), "get_stats")
        return s.as_dict()

    def permutation(self):
        perm = np.empty(self.n, dtype=np.uint64)
        scat = np.empty(self.nnz, dtype=np.uint64)
        self._check(self._lib.b200kkt_get_permutation(self._h, _ptr(perm), _ptr(scat)), "get_permutation")
        return perm.astype(np.int64), scat.astype(np.int64)

    def plan(self):
        sizes = (C.c_int64 * 8)()
        self._check(self._lib.b200kkt_plan_sizes(self._h, sizes), "plan_sizes")
        nsn, nrows = int(sizes[0]), int(sizes[1])
        out = {"num_fronts": nsn, "pool_doubles": int(sizes[2]), "num_levels": int(sizes[3]), "sum_boundary": int(sizes[4]),
               "assembly_distinct": int(sizes[5]), "cb_arena_doubles": int(sizes[6])}
        i32 = lambda: np.empty(nsn, dtype=np.int32)  # noqa: E731
        k, f, ld, parent, level, first = i32(), i32(), i32(), i32(), i32(), i32()
        off = np.empty(nsn, dtype=np.int64)
        self._check(self._lib.b200kkt_plan_fronts(self._h, _ptr(k), _ptr(f), _ptr(ld), _ptr(off), _ptr(parent), _ptr(level), _ptr(first)),
                    "plan_fronts")
        rows_off = np.empty(nsn + 1, dtype=np.int64)
        rows = np.empty(nrows, dtype=np.int32)
        self._check(self._lib.b200kkt_plan_rows(self._h, _ptr(rows_off), _ptr(rows)), "plan_rows")
        out.update(k=k, f=f, ld=ld, pool_offset=off, parent=parent, level=level, first_col=first, rows_off=rows_off, rows=rows)
        return out

    def debug_delay_columns(self, positions):
        p = np.ascontiguousarray(positions, dtype=np.int64)
        self._check(self._lib.b200kkt_debug_delay_columns(self._h, p.size, _ptr(p)), "debug_delay_columns")

    def debug_assemble(self, values, want_scale=True):
        v = np.ascontiguousarray(values, dtype=np.float64)
        pool = np.empty(self.plan()["pool_doubles"], dtype=np.float64)
        scale = np.empty(self.n, dtype=np.float64) if want_scale else None
        self._check(self._lib.b200kkt_debug_assemble(self._h, _ptr(v), _ptr(pool), _ptr(scale)), "debug_assemble")
        return pool, scale

    def debug_factor(self, want_pool=True):
        pool = np.empty(self.plan()["pool_doubles"], dtype=np.float64) if want_pool else None
        dd = np.empty(self.n, dtype=np.float64)
        de = np.empty(self.n, dtype=np.float64)
        pt = np.empty(self.n, dtype=np.int32)
        lp = np.empty(self.n, dtype=np.int32)
        self._check(self._lib.b200kkt_debug_factor(self._h, _ptr(pool), _ptr(dd), _ptr(de), _ptr(pt), _ptr(lp)), "debug_factor")
        return pool, dd, de, pt, lp
