"""In-tree native build: nvcc (sm_100a) for the kernel library, g++ for the host runtime.

Two shared objects are produced under ``distributed_model_parallel_b200/_lib/``:

* ``libdmp_b200.so``     -- every ``csrc/**/*.cu`` (compute + communication kernels, C ABI, see ``include/dmp_b200.h``)
* ``libdmp_b200_rt.so``  -- ``csrc/runtime/*.cpp`` (Reducer bookkeeping, bucket assignment, partitioner, schedules)

The build is incremental (object files are rebuilt only when a source or header is newer) and cross-compiles
without a GPU.  ``python -m distributed_model_parallel_b200._build`` runs it from the command line.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
CSRC = PKG / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = PKG / "_lib"
OBJDIR = LIBDIR / "obj"

KERNEL_LIB = LIBDIR / "libdmp_b200.so"
RUNTIME_LIB = LIBDIR / "libdmp_b200_rt.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC,-fvisibility=hidden",
    "-Xptxas", "-v",
]


def _nvcc() -> str | None:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if cand and os.path.exists(cand) else None


def _newer(src_list, target: Path) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in src_list)


def _headers():
    return list(CSRC.rglob("*.cuh")) + list(CSRC.rglob("*.h")) + list(INCLUDE.glob("*.h"))


def _run(cmd, log: Path | None = None):
    p = subprocess.run(cmd, capture_output=True, text=True)
    if log is not None:
        log.write_text(p.stdout + p.stderr)
    if p.returncode != 0:
        raise RuntimeError("build command failed:\n  %s\n%s\n%s" % (" ".join(map(str, cmd)), p.stdout, p.stderr))
    return p


def build_kernels(force: bool = False, verbose: bool = False) -> Path:
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found: cannot build libdmp_b200.so")
    OBJDIR.mkdir(parents=True, exist_ok=True)
    sources = sorted(p for p in CSRC.rglob("*.cu"))
    headers = _headers()
    jobs = []
    objs = []
    for src in sources:
        obj = OBJDIR / (src.relative_to(CSRC).as_posix().replace("/", "_") + ".o")
        objs.append(obj)
        if force or _newer([src] + headers, obj):
            cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
            jobs.append((cmd, obj.with_suffix(".log")))
    if jobs:
        if verbose:
            print("[dmp_b200] compiling %d CUDA sources for sm_100a" % len(jobs), file=sys.stderr)
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda j: _run(*j), jobs))
    if force or jobs or _newer(objs, KERNEL_LIB):
        # shared CUDA runtime: inside a PyTorch process the library then uses the very libcudart torch has loaded (one runtime
        # instance, one view of current device / peer access); standalone users get it from the toolkit via the rpath
        _run([nvcc, "-shared", "-cudart", "shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xlinker", "-rpath,/usr/local/cuda/lib64",
              "-o", str(KERNEL_LIB), *map(str, objs)])
    return KERNEL_LIB


def build_runtime(force: bool = False, verbose: bool = False) -> Path:
    cxx = os.environ.get("CXX") or shutil.which("g++")
    if cxx is None:
        raise RuntimeError("g++ not found: cannot build libdmp_b200_rt.so")
    LIBDIR.mkdir(parents=True, exist_ok=True)
    sources = sorted((CSRC / "runtime").glob("*.cpp"))
    if force or _newer(list(sources) + _headers(), RUNTIME_LIB):
        if verbose:
            print("[dmp_b200] compiling host runtime", file=sys.stderr)
        _run([cxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden", "-Wall", "-I", str(INCLUDE),
              "-o", str(RUNTIME_LIB), *map(str, sources), "-lpthread"])
    return RUNTIME_LIB


def build_all(force: bool = False, verbose: bool = True):
    return build_kernels(force, verbose), build_runtime(force, verbose)


def ptxas_report() -> str:
    """Concatenated `-Xptxas -v` output (registers / spills / smem) of the last build."""
    out = []
    for log in sorted(OBJDIR.glob("*.log")):
        out.append("== %s ==\n%s" % (log.name, log.read_text()))
    return "\n".join(out)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv)
    if "--report" in sys.argv:
        print(ptxas_report())
    print(KERNEL_LIB, RUNTIME_LIB)
