"""CPU arm of bench.py (TEST INFRASTRUCTURE: only bench.py's cpu_baseline / --impl reference legs import this).

Times the reference's path - numeric factorization + inertia + solve, MUMPSSolver.cpp:85-96,124-147 - on the host cores with the
CPU oracle (oracle/kkt_oracle.cpp) in its restricted-pivoting mode, on the SAME ordering the GPU run uses.  Everything reported
is measured on the workload it names; nothing is projected into `value`.

The ordering comes from a child process (the product's host-only symbolic phase writes the permutation to a file), so the
process that is timed never maps the product's library.  MUMPS itself: tools/mumps_harness.c is built and run by
`mumps_probe()` when a libdmumps is found on the box; otherwise that is recorded as a plain fact.
"""
from __future__ import annotations

import ctypes.util
import glob
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent

_ORDERING_CHILD = r"""
import sys, json
import numpy as np
sys.path.insert(0, {root!r})
import bench
from uno_b200 import KKTSolver, generators as G
kkt = bench.WORKLOADS[{name!r}][{slot}](G, {seed})
s = KKTSolver(device=-1)
s.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, kkt.reg_size, kkt.n, values=kkt.val)
perm, _ = s.permutation()
st = s.stats()
np.save({out!r}, perm)
print(json.dumps({{"flops": st["flops_factor"], "nnz_factor": st["nnz_factor"], "pool_doubles": st["pool_doubles"], "max_front": st["max_front"],
                  "cb_arena_doubles": s.plan()["cb_arena_doubles"], "N": kkt.dim}}))
"""


def ordering_from_child(name: str, slot: int, seed: int = 0):
    """METIS ordering on the KKT-pair-compressed graph (the GPU run's ordering), computed by a child process."""
    import json
    with tempfile.TemporaryDirectory() as tmp:
        out = str(Path(tmp) / "perm.npy")
        code = _ORDERING_CHILD.format(root=str(ROOT), name=name, slot=slot, seed=seed, out=out)
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("ordering child failed: " + r.stderr[-2000:])
        info = json.loads(r.stdout.strip().splitlines()[-1])
        return np.load(out), info


def host_memory_available_gb() -> float:
    try:
        for line in Path("/proc/meminfo").read_text().splitlines():
            if line.startswith("MemAvailable:"):
                return float(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def peak_rss_gb() -> float:
    """Resident-set high-water mark of this process (the oracle's factors, contribution blocks and workspaces live in it)."""
    try:
        for line in Path("/proc/self/status").read_text().splitlines():
            if line.startswith("VmHWM:"):
                return float(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def oracle_memory_estimate_gb(info) -> float:
    """Host memory of one oracle factorization, from the plan of the same ordering: the packed factors, the dense workspace of the
    largest front, the contribution blocks alive at the peak (the oracle sweeps the top of the tree in postorder, which keeps fewer
    alive than the level-order arena of the GPU plan), the generator's and the analysis' own arrays.  Checked against the resident
    peak of the 64^3 boundary-control sample: 7.8 GB estimated, 7.5 GB measured."""
    return 8e-9 * (info["nnz_factor"] + 1.07 * info["max_front"] ** 2 + 0.8 * info["cb_arena_doubles"]) + 1.5e-6 * info.get("N", 0) + 0.5


def run_oracle(kkt, perm, steps: int, warmup: int, threads: int):
    """factorize + inertia + solve per step, wall clock.  Returns (seconds per step, list of times, inertia, backward error, stats)."""
    from oracle import oracle as O
    has_blas = O.bind_openblas()
    o = O.OracleSolver(threshold_u=0.01, threads=threads, restricted=True)
    try:
        t0 = time.perf_counter()
        o.do_symbolic_analysis(kkt.dim, kkt.row, kkt.col, perm)
        t_analyse = time.perf_counter() - t0
        times, t_factor = [], []
        x = None
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            o.do_numerical_factorization(kkt.val)      # raises when the host is about to run out of memory (the oracle watches it)
            t1 = time.perf_counter()
            ine = o.get_inertia()
            x = o.solve_indefinite_system(kkt.rhs)
            t2 = time.perf_counter()
            if it >= warmup:
                times.append(t2 - t0)
                t_factor.append(t1 - t0)
        st = o.stats()
    finally:
        o.close()                                      # factors, workspaces and pools go back to the host
    A = kkt.to_scipy()
    r = kkt.rhs - A @ x
    berr = float(np.abs(r).max() / (abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(kkt.rhs).max()))
    return {"seconds_per_step": float(np.mean(times)), "factor_seconds": float(np.mean(t_factor)), "times": times, "inertia": list(ine),
            "backward_error_first_sweep": berr, "delayed_pivots": int(st["delayed"]), "openblas": bool(has_blas), "analyse_seconds": t_analyse}


def host_dgemm_peak_gflops(n: int = 4096, repeats: int = 3) -> float:
    """DGEMM n^3 on all host cores through numpy's BLAS, best of `repeats`: the ceiling of ANY CPU factorization on this box."""
    rng = np.random.default_rng(0)
    a = rng.standard_normal((n, n))
    b = rng.standard_normal((n, n))
    a @ b
    best = 1e300
    for _ in range(repeats):
        t0 = time.perf_counter()
        a @ b
        best = min(best, time.perf_counter() - t0)
    return 2.0 * n ** 3 / best / 1e9


def cpu_model() -> str:
    try:
        for line in Path("/proc/cpuinfo").read_text().splitlines():
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def mumps_probe():
    """The reference's own arithmetic is MUMPS (un-vendored Fortran).  If the box carries a libdmumps and its header, build and
    run tools/mumps_harness.c (control block of MUMPSSolver.cpp:16-37,82 + ICNTL(7)=1/PERM_IN); else say so."""
    lib = ctypes.util.find_library("dmumps") or ctypes.util.find_library("dmumps_seq")
    hdr = [p for pat in ("/usr/include/dmumps_c.h", "/usr/include/*/dmumps_c.h", "/usr/local/include/dmumps_c.h", "/opt/*/include/dmumps_c.h")
           for p in glob.glob(pat)]
    if not lib or not hdr:
        return {"available": False, "note": "MUMPS not installed on this box (no libdmumps / dmumps_c.h): tools/mumps_harness.c not built"}
    exe = ROOT / "oracle" / "_build" / "mumps_harness"
    exe.parent.mkdir(exist_ok=True)
    cmd = ["gcc", "-O2", "-fopenmp", f"-I{Path(hdr[0]).parent}", str(ROOT / "tools" / "mumps_harness.c"), "-o", str(exe), f"-l{Path(lib).stem[3:].split('.')[0]}", "-lm"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        return {"available": False, "note": "libdmumps found but tools/mumps_harness.c did not link: " + r.stderr[-300:]}
    return {"available": True, "exe": str(exe)}


def run_mumps(exe: str, kkt, perm, threads: int):
    """Hand the matrix, the right-hand side and the ordering to the harness through a binary file; parse its one JSON line."""
    import json
    with tempfile.TemporaryDirectory() as tmp:
        path = Path(tmp) / "system.bin"
        with open(path, "wb") as f:
            np.array([kkt.dim, kkt.nnz], dtype=np.int64).tofile(f)
            (kkt.row.astype(np.int32) + 1).tofile(f)
            (kkt.col.astype(np.int32) + 1).tofile(f)
            kkt.val.astype(np.float64).tofile(f)
            kkt.rhs.astype(np.float64).tofile(f)
            # PERM_IN(i) = position of variable i in the elimination order, 1-based
            pos = np.empty(kkt.dim, dtype=np.int32)
            pos[np.asarray(perm)] = np.arange(1, kkt.dim + 1, dtype=np.int32)
            pos.tofile(f)
        env = dict(os.environ, OMP_NUM_THREADS=str(threads))
        r = subprocess.run([exe, str(path)], capture_output=True, text=True, env=env)
        if r.returncode != 0:
            return {"error": r.stderr[-500:]}
        return json.loads(r.stdout.strip().splitlines()[-1])
