"""In-tree build of the native libraries.

Two shared objects are produced under ``distributed_b200/_lib``:

* ``libb200dist_cuda.so`` — every ``csrc/cuda/*.cu`` and ``csrc/runtime/*.cpp`` compiled by
  ``nvcc`` for ``sm_100a`` only (``-gencode arch=compute_100a,code=sm_100a -lineinfo``).
  nvcc cross-compiles without a GPU, so this runs on the CPU-only build host.
* ``libb200dist_cpu.so`` — ``csrc/cpu/*.cpp`` compiled by ``g++`` (C++ reference ops and the
  shared-memory all-reduce used as numerical oracle and CPU backend).

The build is incremental (object files are rebuilt when their source or any header is
newer) and parallel.  The ``.so`` files are git-ignored but travel to the GPU box with the
repository snapshot.
"""
from __future__ import annotations

import concurrent.futures as cf
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIB_DIR = ROOT / "distributed_b200" / "_lib"
OBJ_DIR = ROOT / "build" / "obj"

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = NVCC_ARCH + [
    "-lineinfo", "-O3", "-std=c++17", "--use_fast_math", "-Xcompiler", "-fPIC",
    "-Xcompiler", "-Wall", f"-I{INCLUDE}",
    "-DB2D_BUILDING=1",
]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-pthread", "-Wall",
             "-march=x86-64-v3", f"-I{INCLUDE}"]


def _nvcc() -> str:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(cand).exists():
        raise RuntimeError("nvcc not found; set NVCC or add /usr/local/cuda/bin to PATH")
    return cand


def _newest_header_mtime() -> float:
    hs = list(INCLUDE.glob("*.h")) + list(CSRC.rglob("*.h")) + list(CSRC.rglob("*.cuh"))
    return max((h.stat().st_mtime for h in hs), default=0.0)


def _stale(src: Path, obj: Path, hdr_mtime: float) -> bool:
    if not obj.exists():
        return True
    t = obj.stat().st_mtime
    return src.stat().st_mtime > t or hdr_mtime > t


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("command failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    # ptxas warnings are worth seeing but not fatal
    if r.stderr.strip() and os.environ.get("B2D_BUILD_VERBOSE"):
        sys.stderr.write(r.stderr)


def _compile_all(jobs: list[tuple[list[str], Path]]) -> None:
    if not jobs:
        return
    with cf.ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
        futs = [ex.submit(_run, cmd) for cmd, _ in jobs]
        for f in futs:
            f.result()


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    """Compile every CUDA / runtime source for sm_100a and link libb200dist_cuda.so."""
    nvcc = _nvcc()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    hdr = _newest_header_mtime()
    srcs = sorted((CSRC / "cuda").glob("*.cu")) + sorted((CSRC / "runtime").glob("*.cpp"))
    jobs, objs = [], []
    for s in srcs:
        o = OBJ_DIR / ("cuda_" + s.stem + ".o")
        objs.append(o)
        if force or _stale(s, o, hdr):
            extra = ["-x", "cu"] if s.suffix == ".cpp" else []
            jobs.append(([nvcc, *NVCC_FLAGS, *extra, "-c", str(s), "-o", str(o)], o))
    if verbose and jobs:
        print(f"[b200dist] nvcc: compiling {len(jobs)} file(s) for sm_100a", file=sys.stderr)
    _compile_all(jobs)
    lib = LIB_DIR / "libb200dist_cuda.so"
    if jobs or not lib.exists():
        _run([nvcc, *NVCC_ARCH, "-shared", "-o", str(lib), *map(str, objs), "-ldl", "-lpthread"])
    return lib


def build_cpu(force: bool = False, verbose: bool = False) -> Path:
    """Compile the C++ CPU backend and link libb200dist_cpu.so."""
    cxx = os.environ.get("CXX") or shutil.which("g++") or "g++"
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    hdr = _newest_header_mtime()
    srcs = sorted((CSRC / "cpu").glob("*.cpp"))
    jobs, objs = [], []
    for s in srcs:
        o = OBJ_DIR / ("cpu_" + s.stem + ".o")
        objs.append(o)
        if force or _stale(s, o, hdr):
            jobs.append(([cxx, *CXX_FLAGS, "-c", str(s), "-o", str(o)], o))
    if verbose and jobs:
        print(f"[b200dist] g++: compiling {len(jobs)} file(s)", file=sys.stderr)
    _compile_all(jobs)
    lib = LIB_DIR / "libb200dist_cpu.so"
    if jobs or not lib.exists():
        _run([cxx, "-shared", "-pthread", "-o", str(lib), *map(str, objs), "-lrt", "-lpthread"])
    return lib


def build_all(force: bool = False, verbose: bool = True) -> dict[str, Path]:
    return {"cpu": build_cpu(force, verbose), "cuda": build_cuda(force, verbose)}


if __name__ == "__main__":
    out = build_all(force="--force" in sys.argv)
    for k, v in out.items():
        print(f"{k}: {v}")
