"""In-tree build of the native libraries (no pip install, no JIT cache).  The reference ships no
build step at all (/root/reference: simple_distributed.py + readme.md).

Two shared objects are produced under ``simple_distributed_machine_learning_b200/_lib``:

* ``libsdml_b200.so``  -- every hand-written sm_100a CUDA kernel (``csrc/*.cu``), compiled
  with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``; C ABI declared in
  ``include/sdml_kernels.h`` and loaded with :mod:`ctypes`.
* ``libsdml_sched.so`` -- the C++ pipeline-schedule generator / simulator
  (``csrc/schedule.cpp``), plain g++.

The built ``.so`` files are git-ignored but travel to the GPU box with the gpurun
snapshot.  ``build()`` is idempotent: it only recompiles when a source is newer
than the library.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent.parent
REPO_DIR = PKG_DIR.parent
CSRC = REPO_DIR / "csrc"
INCLUDE = REPO_DIR / "include"
LIB_DIR = PKG_DIR / "_lib"

CUDA_LIB = LIB_DIR / "libsdml_b200.so"
SCHED_LIB = LIB_DIR / "libsdml_sched.so"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "--compiler-options", "-fPIC",
    "-Xptxas", "-v",
]


def _nvcc() -> str | None:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else None


def _stale(target: Path, sources: list[Path]) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(s.stat().st_mtime > t for s in sources)


def _run(cmd: list[str], log: Path) -> None:
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    log.write_text(" ".join(cmd) + "\n" + proc.stdout)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout[-8000:])
        raise RuntimeError(f"build failed: {' '.join(cmd[:3])} ... (log: {log})")


class _BuildLock:
    """Cross-process lock around a rebuild: `launch.py` starts N ranks at once, and with a stale
    library they would all run nvcc into the same object files while another rank dlopen()s a
    half-written .so.  One rank builds, the others wait and then find the library fresh."""

    def __enter__(self):
        import fcntl
        LIB_DIR.mkdir(parents=True, exist_ok=True)
        self._f = open(LIB_DIR / ".build.lock", "w")
        fcntl.flock(self._f, fcntl.LOCK_EX)
        return self

    def __exit__(self, *exc):
        import fcntl
        fcntl.flock(self._f, fcntl.LOCK_UN)
        self._f.close()


def cuda_sources_newer() -> bool:
    """True when a CUDA source / header is newer than the built library (or it is missing)."""
    cu_sources = sorted(CSRC.glob("*.cu"))
    headers = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted(INCLUDE.glob("*.h"))
    return _stale(CUDA_LIB, cu_sources + headers)


def build_cuda(force: bool = False, verbose: bool = False) -> Path:
    """Compile every ``csrc/*.cu`` into ``libsdml_b200.so`` (one object per source, in parallel)."""
    if not force and not cuda_sources_newer():
        return CUDA_LIB
    with _BuildLock():
        return _build_cuda_locked(force, verbose)


def _build_cuda_locked(force: bool, verbose: bool) -> Path:
    cu_sources = sorted(CSRC.glob("*.cu"))
    headers = sorted(CSRC.glob("*.cuh")) + sorted(CSRC.glob("*.h")) + sorted(INCLUDE.glob("*.h"))
    if not force and not _stale(CUDA_LIB, cu_sources + headers):
        return CUDA_LIB          # another process built it while we waited for the lock
    nvcc = _nvcc()
    if nvcc is None:
        raise RuntimeError("nvcc not found; cannot build libsdml_b200.so")
    obj_dir = REPO_DIR / "build" / "obj"
    obj_dir.mkdir(parents=True, exist_ok=True)
    procs = []
    objs = []
    for src in cu_sources:
        obj = obj_dir / (src.stem + ".o")
        objs.append(obj)
        if not force and not _stale(obj, [src] + headers):
            continue
        cmd = [nvcc, *NVCC_FLAGS, "-I", str(INCLUDE), "-I", str(CSRC), "-c", str(src), "-o", str(obj)]
        log = obj_dir / (src.stem + ".log")
        procs.append((src, log, cmd, subprocess.Popen(cmd, stdout=open(log, "w"), stderr=subprocess.STDOUT)))
    failed = []
    for src, log, cmd, p in procs:
        if p.wait() != 0:
            failed.append((src, log))
        elif verbose:
            sys.stderr.write(log.read_text())
    if failed:
        for src, log in failed:
            sys.stderr.write(f"--- {src} ---\n{log.read_text()[-6000:]}\n")
        raise RuntimeError(f"nvcc failed for {[str(s) for s, _ in failed]}")
    tmp_lib = CUDA_LIB.with_suffix(f".so.tmp{os.getpid()}")
    link = [nvcc, "-shared", "-o", str(tmp_lib), *map(str, objs), "-lcudart"]
    _run(link, obj_dir / "link.log")
    os.replace(tmp_lib, CUDA_LIB)      # readers see the old or the new library, never half of one
    return CUDA_LIB


def build_sched(force: bool = False) -> Path:
    """Compile the C++ schedule generator into ``libsdml_sched.so``."""
    LIB_DIR.mkdir(parents=True, exist_ok=True)
    sources = sorted(CSRC.glob("*.cpp"))
    headers = sorted(INCLUDE.glob("*.h"))
    if not force and not _stale(SCHED_LIB, sources + headers):
        return SCHED_LIB
    with _BuildLock():
        if not force and not _stale(SCHED_LIB, sources + headers):
            return SCHED_LIB
        cxx = shutil.which("g++") or "g++"
        obj_dir = REPO_DIR / "build" / "obj"
        obj_dir.mkdir(parents=True, exist_ok=True)
        tmp_lib = SCHED_LIB.with_suffix(f".so.tmp{os.getpid()}")
        cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-I", str(INCLUDE),
               "-o", str(tmp_lib), *map(str, sources)]
        _run(cmd, obj_dir / "sched.log")
        os.replace(tmp_lib, SCHED_LIB)
    return SCHED_LIB


def build_all(force: bool = False, verbose: bool = False) -> None:
    build_sched(force)
    build_cuda(force, verbose)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(CUDA_LIB, SCHED_LIB)
