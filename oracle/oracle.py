"""ctypes wrapper of the CPU oracle (oracle/kkt_oracle.cpp).  TEST INFRASTRUCTURE ONLY: imported by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by uno_b200."""
from __future__ import annotations

import ctypes as C
import glob
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "_build" / "libkkt_oracle.so"
SRC = HERE / "kkt_oracle.cpp"
METIS = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda")) / "targets/x86_64-linux/lib/libmetis_static.a"


def build(force: bool = False) -> Path:
    if LIB.exists() and not force and LIB.stat().st_mtime >= SRC.stat().st_mtime:
        return LIB
    LIB.parent.mkdir(exist_ok=True)
    cmd = ["g++", "-std=c++17", "-O3", "-mavx2", "-mfma", "-fopenmp", "-fPIC", "-shared", str(SRC), str(METIS), "-ldl", "-lm", "-o", str(LIB)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stderr)
    return LIB


_lib = None


def _load():
    global _lib
    if _lib is None:
        # workers spin for about a millisecond after a parallel region, then sleep (the pivot loop opens short regions back to back)
        os.environ.setdefault("GOMP_SPINCOUNT", "3000000")
        if not LIB.exists():
            build()
        lib = C.CDLL(str(LIB))
        lib.oracle_create.restype = C.c_void_p
        lib.oracle_destroy.argtypes = [C.c_void_p]
        lib.oracle_last_error.argtypes = [C.c_void_p]
        lib.oracle_last_error.restype = C.c_char_p
        lib.oracle_set_blas.argtypes = [C.c_char_p, C.c_char_p]
        lib.oracle_analyse.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_set_parameters.argtypes = [C.c_void_p, C.c_double, C.c_double]
        lib.oracle_set_pivot_mode.argtypes = [C.c_void_p, C.c_int]
        lib.oracle_factorize.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.oracle_inertia.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        lib.oracle_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        lib.oracle_get_permutation.argtypes = [C.c_void_p, C.c_void_p]
        lib.oracle_stats.argtypes = [C.c_void_p, C.c_void_p]
        _lib = lib
    return _lib


def bind_openblas() -> bool:
    """Use the OpenBLAS that ships inside scipy for the oracle's dense updates (threaded DGEMM),
    as MUMPS would use a threaded BLAS.  Falls back to plain loops when it cannot be found."""
    lib = _load()
    try:
        import scipy
        cands = glob.glob(str(Path(scipy.__file__).resolve().parent.parent / "scipy.libs" / "libscipy_openblas*.so"))
    except Exception:
        cands = []
    for path in cands:
        for prefix in (b"scipy_", b""):
            if lib.oracle_set_blas(path.encode(), prefix):
                return True
    return False


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleSolver:
    """CPU multifrontal LDL^T with threshold pivoting and delayed pivots (see kkt_oracle.cpp)."""

    def __init__(self, threshold_u: float = 0.01, null_tol: float = 2.2e-21, threads: int = 1, restricted: bool = False):
        """restricted = False: threshold test against the whole front column (MUMPS' rule); True: against the fully-summed
        rows only (few delayed pivots: the mode for BASELINE-size checks and for the same-flops CPU baseline)."""
        self._lib = _load()
        self._h = C.c_void_p(self._lib.oracle_create())
        self._lib.oracle_set_parameters(self._h, threshold_u, null_tol)
        self._lib.oracle_set_pivot_mode(self._h, 1 if restricted else 0)
        self.threads = threads
        self.n = 0

    def close(self):
        if getattr(self, "_h", None):
            self._lib.oracle_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError("oracle: " + self._lib.oracle_last_error(self._h).decode())

    def do_symbolic_analysis(self, n, row, col, perm=None):
        row = np.ascontiguousarray(row, dtype=np.uint64)
        col = np.ascontiguousarray(col, dtype=np.uint64)
        perm = None if perm is None else np.ascontiguousarray(perm, dtype=np.int64)
        self.n, self.nnz = int(n), int(row.size)
        self._check(self._lib.oracle_analyse(self._h, self.n, self.nnz, _p(row), _p(col), _p(perm)))

    def do_numerical_factorization(self, values):
        v = np.ascontiguousarray(values, dtype=np.float64)
        assert v.size == self.nnz
        self._check(self._lib.oracle_factorize(self._h, _p(v), self.threads))

    def get_inertia(self):
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._lib.oracle_inertia(self._h, C.byref(a), C.byref(b), C.byref(c))
        return int(a.value), int(b.value), int(c.value)

    def matrix_is_singular(self):
        return self.get_inertia()[2] > 0

    def solve_indefinite_system(self, rhs):
        b = np.ascontiguousarray(rhs, dtype=np.float64)
        x = np.empty(self.n)
        self._check(self._lib.oracle_solve(self._h, _p(b), _p(x)))
        return x

    def permutation(self):
        p = np.empty(self.n, dtype=np.int64)
        self._lib.oracle_get_permutation(self._h, _p(p))
        return p

    def stats(self):
        out = np.zeros(10)
        self._lib.oracle_stats(self._h, _p(out))
        keys = ["nnz_factor", "flops", "delayed", "two_by_two", "t_analyse", "t_factor", "t_solve", "fronts", "blas", "omp_threads"]
        return dict(zip(keys, out.tolist()))
