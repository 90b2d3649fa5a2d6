"""Recipe (test infrastructure) that compiles the UNCHANGED Uno host code from the read-only reference tree and
links it with this repo's Uno adapter into oracle/_ref/uno_driver.

* Reference sources are compiled where they lie under /root/reference (never copied); objects and the binary go
  to oracle/_ref/ (git-ignored, shipped to the GPU box with the snapshot).
* The one reference file that is NOT compiled is uno/ingredients/subproblem_solvers/SymmetricIndefiniteLinearSolverFactory.cpp:
  uno_b200/uno_plugin/SymmetricIndefiniteLinearSolverFactory.cpp replaces it (registers "B200", and "ORACLE" for
  the reference runs).  Source list = the GLOB of the reference's CMakeLists.txt:62-83.
* The driver links libb200kkt.so (product) and the CPU oracle (stand-in for MUMPS, which cannot be built here).
The reference's own build system is not used.  On the GPU box /root/reference does not exist: the prebuilt binary is used.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
REF = Path(os.environ.get("UNO_REFERENCE", "/root/reference"))
OUT = ROOT / "oracle" / "_ref"
OBJ = OUT / "obj"
PLUGIN = ROOT / "uno_b200" / "uno_plugin"
BINARY = OUT / "uno_driver"
CUDA_HOME = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))
METIS = CUDA_HOME / "targets/x86_64-linux/lib/libmetis_static.a"

REFERENCE_GLOBS = [
    "uno/Uno.cpp",
    "uno/ingredients/constraint_relaxation_strategies/*.cpp",
    "uno/ingredients/globalization_mechanisms/*.cpp",
    "uno/ingredients/globalization_strategies/*.cpp",
    "uno/ingredients/globalization_strategies/switching_methods/*.cpp",
    "uno/ingredients/globalization_strategies/switching_methods/filter_methods/*.cpp",
    "uno/ingredients/globalization_strategies/switching_methods/filter_methods/filters/*.cpp",
    "uno/ingredients/globalization_strategies/switching_methods/funnel_methods/*.cpp",
    "uno/ingredients/hessian_models/*.cpp",
    "uno/ingredients/inequality_handling_methods/*.cpp",
    "uno/ingredients/inequality_handling_methods/inequality_constrained_methods/*.cpp",
    "uno/ingredients/inequality_handling_methods/interior_point_methods/*.cpp",
    "uno/ingredients/regularization_strategies/*.cpp",
    "uno/ingredients/subproblem/*.cpp",
    "uno/ingredients/subproblem_solvers/*.cpp",
    "uno/model/*.cpp",
    "uno/optimization/*.cpp",
    "uno/options/*.cpp",
    "uno/preprocessing/*.cpp",
    "uno/tools/*.cpp",
]
REPLACED = {"SymmetricIndefiniteLinearSolverFactory.cpp"}
CXXFLAGS = ["-std=c++17", "-O2", "-fPIC", "-DNDEBUG", "-DHAS_B200", "-DHAS_KKT_ORACLE"]


def _compile(src: Path, obj: Path, includes, extra=()):
    if obj.exists() and obj.stat().st_mtime >= src.stat().st_mtime:
        return
    cmd = ["g++", *CXXFLAGS, *extra, *[f"-I{i}" for i in includes], "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"compiling {src} failed:\n{r.stderr[-4000:]}")


def available() -> bool:
    return (REF / "uno" / "Uno.cpp").exists()


def build(optional: bool = False, verbose: bool = False) -> Path | None:
    if not available():
        if optional:
            return BINARY if BINARY.exists() else None
        raise RuntimeError(f"reference tree not found at {REF}")
    OBJ.mkdir(parents=True, exist_ok=True)
    includes = [REF / "uno", PLUGIN, ROOT / "include", ROOT / "oracle" / "uno_oracle"]
    jobs = []
    for pattern in REFERENCE_GLOBS:
        for src in sorted(REF.glob(pattern)):
            if src.name in REPLACED:
                continue
            rel = src.relative_to(REF).with_suffix("")
            jobs.append((src, OBJ / ("ref__" + "__".join(rel.parts) + ".o"), ()))
    own = [PLUGIN / "B200Solver.cpp", PLUGIN / "SymmetricIndefiniteLinearSolverFactory.cpp", PLUGIN / "uno_driver.cpp",
           PLUGIN / "models" / "AnalyticModel.cpp", PLUGIN / "models" / "NLModel.cpp", ROOT / "oracle" / "uno_oracle" / "OracleSolver.cpp"]
    for src in own:
        jobs.append((src, OBJ / ("own__" + src.stem + ".o"), ()))
    jobs.append((ROOT / "oracle" / "kkt_oracle.cpp", OBJ / "own__kkt_oracle.o", ("-O3", "-mavx2", "-mfma", "-fopenmp")))
    # headers of the plugin: rebuild own objects when they change
    newest_header = max(p.stat().st_mtime for p in list(PLUGIN.rglob("*.hpp")) + [ROOT / "include" / "b200kkt.h"])
    for src, obj, _ in jobs:
        if obj.name.startswith("own__") and obj.exists() and obj.stat().st_mtime < newest_header:
            obj.unlink()
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as pool:
        list(pool.map(lambda j: _compile(j[0], j[1], includes, j[2]), jobs))
    objs = [str(j[1]) for j in jobs]
    libdir = ROOT / "uno_b200" / "lib"
    if not (libdir / "libb200kkt.so").exists():
        raise RuntimeError("build libb200kkt.so first (python -m uno_b200.build)")
    if not BINARY.exists() or any(Path(o).stat().st_mtime > BINARY.stat().st_mtime for o in objs) or \
            (libdir / "libb200kkt.so").stat().st_mtime > BINARY.stat().st_mtime:
        cmd = ["g++", "-o", str(BINARY), *objs, str(METIS), f"-L{libdir}", "-lb200kkt", "-Wl,-rpath,$ORIGIN/../../uno_b200/lib",
               "-fopenmp", "-ldl", "-lm", "-lpthread"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("linking uno_driver failed:\n" + r.stderr[-4000:])
    if verbose:
        print("built", BINARY, f"({len(jobs)} objects)")
    return BINARY


if __name__ == "__main__":
    build(verbose=True)
