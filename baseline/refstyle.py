"""`bench.py --impl refstyle`: the reference-STYLE baseline (written by this project).

Same per-iteration library-call sequence as CUDACG.cu:269-352 (cuSPARSE CSR SpMV + cuBLAS level-1 in
host pointer mode with its two blocking scalar read-backs) on the benchmark grid.  Single GPU only: the
reference has no multi-GPU path (CUDACG.cu:87).  CGB_REFSTYLE_FAITHFUL=1 additionally re-queries and
re-allocates the SpMV workspace every iteration like CUDACG.cu:273-285.
"""
from __future__ import annotations

import ctypes as C
import json
import os

from cuda_mpi_parallel_b200.utils import build


def _lib():
    build.build_baseline()
    L = C.CDLL(str(build.BASELINE_LIB))
    L.cgb_refstyle_run.restype = C.c_int
    L.cgb_refstyle_run.argtypes = [C.c_int] * 7 + [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    L.cgb_refstyle_last_error.restype = C.c_char_p
    return L


def run(kind: int, grid, warmup: int, steps: int, faithful: bool):
    L = _lib()
    ms, rn, nnz = C.c_double(), C.c_double(), C.c_longlong()
    gx, gy, gz = (list(grid) + [1, 1])[:3]
    rc = L.cgb_refstyle_run(kind, gx, gy, gz, int(faithful), warmup, steps, C.byref(ms), C.byref(rn), C.byref(nnz))
    if rc != 0:
        raise RuntimeError("refstyle baseline: " + L.cgb_refstyle_last_error().decode())
    return ms.value, rn.value, nnz.value


def bench(a, rank, local_rank, world):
    if world != 1:
        if rank == 0:
            print(json.dumps({"impl": "refstyle", "unavailable":
                              "the reference call sequence is single-GPU (cudaSetDevice(0), CUDACG.cu:87)"}))
        return 0
    kinds = {"poisson2d5": 1, "poisson3d7": 2, "stencil3d27": 3}
    defaults = {"poisson3d7": [512, 512, 512], "poisson2d5": [8192, 8192], "stencil3d27": [384, 384, 384]}
    grid = a.grid or defaults[a.problem]
    if len(grid) == 1:
        grid = grid * (2 if a.problem == "poisson2d5" else 3)
    faithful = os.environ.get("CGB_REFSTYLE_FAITHFUL", "0") == "1"
    K, W = a.steps, max(a.warmup, 3)
    ms, rn, nnz = run(kinds[a.problem], grid, W, K, faithful)
    n = 1
    for g in grid:
        n *= g
    its = K / (ms * 1e-3)
    print(json.dumps({
        "metric": f"CG iterations/s ({a.problem} {'x'.join(map(str, grid))} fp64)", "value": its,
        "unit": "CG iterations/s", "n_gpus": 1, "steps": K, "warmup": W, "ms_per_step": ms / K,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp64",
        "data": "synthetic (hash-random RHS, assembled CSR stencil operator)", "gdof_per_s": its * n / 1e9,
        "impl": "refstyle (reference-style cuSPARSE/cuBLAS baseline written by this project; "
                + ("faithful: per-iteration workspace malloc" if faithful else "hoisted workspace") + ")",
        "config": {"model": f"{a.problem} {'x'.join(map(str, grid))} fp64 unpreconditioned CG", "nnz": nnz,
                   "last_rnorm": rn}}))
    return 0
