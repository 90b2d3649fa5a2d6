"""In-tree build of libb200kkt.so (C ABI + sm_100a kernels).  nvcc cross-compiles without a GPU.

The library is the product: host symbolic phase (C++), CUDA numeric phase, C ABI
(include/b200kkt.h).  It is built in-tree (uno_b200/lib/) so it travels with the repo snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
LIBDIR = ROOT / "lib"
LIB = LIBDIR / "libb200kkt.so"
CUDA_HOME = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))
METIS = CUDA_HOME / "targets/x86_64-linux/lib/libmetis_static.a"
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-Wall", "-Wextra", f"-I{CUDA_HOME}/include"]


def _newer(target: Path, sources) -> bool:
    if not target.exists():
        return False
    t = target.stat().st_mtime
    return all(Path(s).stat().st_mtime <= t for s in sources)


def _run(cmd):
    r = subprocess.run([str(c) for c in cmd], capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(str(c) for c in cmd))


def build(force: bool = False, verbose: bool = False) -> Path:
    LIBDIR.mkdir(exist_ok=True)
    headers = sorted(CSRC.glob("*.hpp")) + [ROOT.parent / "include" / "b200kkt.h"]
    objs = []
    for src in sorted(CSRC.glob("*.cu")) + sorted(CSRC.glob("*.cpp")):
        obj = LIBDIR / (src.stem + ".o")
        objs.append(obj)
        if not force and _newer(obj, [src] + headers):
            continue
        if verbose:
            print("compiling", src.name)
        if src.suffix == ".cu":
            _run([CUDA_HOME / "bin/nvcc", *NVCC_FLAGS, "-c", src, "-o", obj])
        else:
            _run(["g++", *CXX_FLAGS, "-c", src, "-o", obj])
    if force or not _newer(LIB, objs):
        if not METIS.exists():
            raise RuntimeError(f"METIS static library not found at {METIS}")
        _run(["g++", "-shared", "-o", LIB, *objs, METIS, f"-L{CUDA_HOME}/lib64", "-lcudart_static", "-ldl", "-lrt", "-lpthread", "-lm"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
