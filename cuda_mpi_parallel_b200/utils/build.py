"""In-tree build of the native library for sm_100a.

The reference's whole build recipe is the comment ``nvcc CUDACG.cu -lcublas -lcusparse``
(CUDACG.cu:1).  Here every translation unit under ``csrc/`` is compiled with
``-gencode arch=compute_100a,code=sm_100a -lineinfo`` and linked into
``cuda_mpi_parallel_b200/lib/libcgb200.so`` (git-ignored, but shipped to the GPU box by
``gpurun``).  ``nvcc`` cross-compiles without a GPU, so this runs on the CPU-only dev box too.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
BUILD = ROOT / "build"
LIBDIR = ROOT / "cuda_mpi_parallel_b200" / "lib"
LIB = LIBDIR / "libcgb200.so"
BASELINE_LIB = LIBDIR / "libcgb200_baseline.so"
CLI = ROOT / "bin" / "cgb200"

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
          f"-I{ROOT / 'include'}", f"-I{CSRC}", "-Xptxas", "-v"]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _sources(sub: Path):
    return sorted([p for p in sub.rglob("*") if p.suffix in (".cu", ".cpp")])


def _headers():
    return sorted(list(CSRC.rglob("*.hpp")) + list(CSRC.rglob("*.cuh")) + list((ROOT / "include").glob("*.h")))


def _digest(paths, extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in paths:
        h.update(str(p).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _compile(src: Path, obj: Path, log: Path, extra=()):
    cmd = [nvcc(), *ARCH, *COMMON, *extra, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log.write_text("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}{r.stderr}")
    return obj


def _build_lib(name: str, srcs, out: Path, link_extra=(), force=False, verbose=False) -> bool:
    """Compile srcs -> objects -> shared library. Returns True if anything was rebuilt."""
    BUILD.mkdir(exist_ok=True)
    LIBDIR.mkdir(exist_ok=True)
    hdr_digest = _digest(_headers(), " ".join(ARCH + COMMON))
    jobs = []
    objs = []
    for src in srcs:
        tag = hashlib.sha1(str(src.relative_to(ROOT)).encode()).hexdigest()[:10]
        obj = BUILD / f"{name}_{src.stem}_{tag}.o"
        stamp = obj.with_suffix(".stamp")
        want = _digest([src], hdr_digest)
        objs.append(obj)
        if force or not obj.exists() or not stamp.exists() or stamp.read_text() != want:
            jobs.append((src, obj, stamp, want))
    rebuilt = bool(jobs)
    if jobs:
        if verbose:
            print(f"[build] compiling {len(jobs)} file(s) for sm_100a", file=sys.stderr)
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            futs = {ex.submit(_compile, s, o, o.with_suffix(".log")): (s, o, st, w) for s, o, st, w in jobs}
            for f in cf.as_completed(futs):
                s, o, st, w = futs[f]
                f.result()
                st.write_text(w)
    if rebuilt or not out.exists():
        cmd = [nvcc(), *ARCH, "-shared", "-o", str(out), *map(str, objs), *link_extra]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}{r.stderr}")
        rebuilt = True
    return rebuilt


def build_core(force=False, verbose=False) -> Path:
    import fcntl
    BUILD.mkdir(exist_ok=True)
    with open(BUILD / ".lock", "w") as lk:      # several ranks may import at once
        fcntl.flock(lk, fcntl.LOCK_EX)
        _build_lib("core", _sources(CSRC), LIB, force=force, verbose=verbose)
    return LIB


B4_EXE = ROOT / "bin" / "refstyle_b4"


def build_baseline(force=False, verbose=False) -> Path:
    srcs = _sources(ROOT / "baseline" / "src") if (ROOT / "baseline" / "src").exists() else []
    srcs = [s for s in srcs if not s.stem.endswith("_main")]
    if not srcs:
        return BASELINE_LIB
    _build_lib("baseline", srcs, BASELINE_LIB, link_extra=["-lcusparse", "-lcublas"], force=force, verbose=verbose)
    return BASELINE_LIB


def build_b4(force=False, verbose=False) -> Path:
    """The multi-GPU reference-style baseline B4 (cuSPARSE/cuBLAS + NCCL all-reduce + peer-copy halo) as a separate
    executable: it links the SYSTEM libnccl, which must not share a process with the NCCL that torch bundles."""
    src = ROOT / "baseline" / "src" / "ref_b4_main.cu"
    if not src.exists():
        return B4_EXE
    B4_EXE.parent.mkdir(exist_ok=True)
    newest = max(p.stat().st_mtime for p in [src, LIB, ROOT / "include" / "cgb200.h"])
    if force or not B4_EXE.exists() or B4_EXE.stat().st_mtime < newest:
        cmd = [nvcc(), *ARCH, "-O3", "-std=c++17", "-lineinfo", f"-I{ROOT / 'include'}", str(src), "-o", str(B4_EXE),
               f"-L{LIBDIR}", "-lcgb200", "-lnccl", "-lcusparse", "-lcublas", "-lpthread",
               "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../cuda_mpi_parallel_b200/lib"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"B4 baseline build failed:\n{r.stdout}{r.stderr}")
    return B4_EXE


def build_cli(force=False, verbose=False) -> Path:
    """The command-line front end: plain C++ on top of the C ABI (no CUDA in this translation unit)."""
    src = ROOT / "tools" / "cgb200_main.cpp"
    if not src.exists():
        return CLI
    CLI.parent.mkdir(exist_ok=True)
    newest = max(p.stat().st_mtime for p in [src, LIB, ROOT / "include" / "cgb200.h"])
    if force or not CLI.exists() or CLI.stat().st_mtime < newest:
        cmd = ["g++", "-O2", "-std=c++17", "-Wall", f"-I{ROOT / 'include'}", str(src), "-o", str(CLI),
               f"-L{LIBDIR}", "-lcgb200", "-Wl,-rpath,$ORIGIN/../cuda_mpi_parallel_b200/lib"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"CLI build failed:\n{r.stdout}{r.stderr}")
    return CLI


def build_all(force=False, verbose=False):
    build_core(force, verbose)
    build_baseline(force, verbose)
    build_b4(force, verbose)
    build_cli(force, verbose)
    return LIB


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose=True)
    print(LIB)
