"""`bench.py --impl refstyle`: the reference-STYLE baselines (written by this project, never upstream's numbers).

B3 (1 GPU): the per-iteration library-call sequence of CUDACG.cu:269-352 (cuSPARSE CSR SpMV + cuBLAS level-1 in host
pointer mode with its two blocking scalar read-backs) on the benchmark grid, in `libcgb200_baseline.so`.
CGB_REFSTYLE_FAITHFUL=1 additionally re-queries and re-allocates the SpMV workspace every iteration like
CUDACG.cu:273-285.

B4 (N > 1 GPUs): the same loop ported the obvious way - z-slab row blocks, ncclAllReduce for the two dot products,
cudaMemcpyPeerAsync + a blocking rendezvous for the halo - in the separate executable `bin/refstyle_b4` (it links the
system NCCL, which must not share a process with torch's). The reference itself has no multi-GPU path (CUDACG.cu:87).

Both use the framework's hash-random right-hand side, so their residuals are comparable with the solver's.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess

from cuda_mpi_parallel_b200.utils import build
from cuda_mpi_parallel_b200.utils.clocks import ClockSampler

KINDS = {"poisson2d5": 1, "poisson3d7": 2, "stencil3d27": 3, "csr7": 2}
DEFAULTS = {"poisson3d7": [512, 512, 512], "poisson2d5": [8192, 8192], "stencil3d27": [384, 384, 384],
            "csr7": [512, 512, 512]}


def _lib():
    build.build_baseline()
    L = C.CDLL(str(build.BASELINE_LIB))
    L.cgb_refstyle_run2.restype = C.c_int
    L.cgb_refstyle_run2.argtypes = [C.c_int] * 7 + [C.c_void_p, C.c_void_p, C.POINTER(C.c_double),
                                                    C.POINTER(C.c_double), C.POINTER(C.c_longlong)]
    L.cgb_refstyle_last_error.restype = C.c_char_p
    return L


def run(kind: int, grid, warmup: int, steps: int, faithful: bool, b_dev=None, hist=None):
    """One B3 run. b_dev: CUDA float64 tensor with the global RHS (None: built-in RHS); hist: numpy float64 array of
    warmup+steps entries that receives ||r||_2 after every iteration. Returns (ms for the timed steps, last ||r||, nnz)."""
    L = _lib()
    ms, rn, nnz = C.c_double(), C.c_double(), C.c_longlong()
    gx, gy, gz = (list(grid) + [1, 1])[:3]
    bp = C.c_void_p(b_dev.data_ptr()) if b_dev is not None else None
    hp = C.c_void_p(hist.ctypes.data) if hist is not None else None
    rc = L.cgb_refstyle_run2(kind, gx, gy, gz, int(faithful), warmup, steps, bp, hp, C.byref(ms), C.byref(rn),
                             C.byref(nnz))
    if rc != 0:
        raise RuntimeError("refstyle baseline: " + L.cgb_refstyle_last_error().decode())
    return ms.value, rn.value, nnz.value


def _grid_of(a):
    grid = a.grid or DEFAULTS[a.problem]
    if len(grid) == 1:
        grid = grid * (2 if a.problem == "poisson2d5" else 3)
    return list(grid)


def run_b4(problem: str, grid, gpus: int, warmup: int, steps: int, timeout: int = 900):
    """One B4 run (separate process). Returns the parsed JSON line of bin/refstyle_b4."""
    build.build_b4()
    kind = KINDS[problem]
    g3 = [grid[0], 1, grid[1]] if kind == 1 else list(grid)       # a 2D grid is nx x 1 x ny, split along its rows
    cmd = [str(build.B4_EXE), "--kind", str(kind), "--grid", *map(str, g3), "--gpus", str(gpus), "--steps", str(steps),
           "--warmup", str(warmup)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not lines:
        raise RuntimeError(f"refstyle_b4 failed rc={r.returncode}: {r.stderr[-800:]}")
    return json.loads(lines[-1])


def bench(a, rank, local_rank, world):
    grid = _grid_of(a)
    K, W = a.steps, max(a.warmup, 3)
    n = 1
    for g in grid:
        n *= g
    base = {"unit": "CG iterations/s", "steps": K, "warmup": W, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "fp64",
            "data": "synthetic (the framework's hash-random RHS, assembled CSR stencil operator)",
            "metric": f"CG iterations/s ({a.problem} {'x'.join(map(str, grid))} fp64)"}
    if world > 1:
        if rank != 0:
            return 0                                # rank 0 drives all GPUs through one multi-threaded process
        with ClockSampler(0) as clk:
            out = run_b4(a.problem, grid, world, W, K)
        print(json.dumps({**base, "value": out["value"], "n_gpus": world, "ms_per_step": out["ms_per_step"],
                          "gdof_per_s": out["gdof_per_s"], "clocks": clk.summary(),
                          "impl": "refstyle-b4 (reference call sequence + ncclAllReduce + cudaMemcpyPeerAsync halo, "
                                  "written by this project)",
                          "r0norm": out["r0norm"], "rnorm_after_steps": out["rnorm_after_steps"],
                          "config": {"model": f"{a.problem} {'x'.join(map(str, grid))} fp64 unpreconditioned CG",
                                     "nnz": out["nnz"], "parallelism": f"z-slab row blocks x{world}",
                                     "clocks_note": "sampled on GPU 0 over the whole baseline process (set-up included)"}}))
        return 0
    import numpy as np
    import torch
    from cuda_mpi_parallel_b200 import Problem
    faithful = os.environ.get("CGB_REFSTYLE_FAITHFUL", "0") == "1"
    prob = Problem.from_name("poisson3d7" if a.problem == "csr7" else a.problem, grid)
    b_dev = torch.from_numpy(prob.rhs("random", seed=1)).cuda()     # the RHS of the solver's kernel arm
    hist = np.zeros(W + K)
    with ClockSampler(local_rank) as clk:
        ms, rn, nnz = run(KINDS[a.problem], grid, W, K, faithful, b_dev, hist)
    its = K / (ms * 1e-3)
    print(json.dumps({**base, "value": its, "n_gpus": 1, "ms_per_step": ms / K, "gdof_per_s": its * n / 1e9,
                      "clocks": clk.summary(),
                      "impl": "refstyle (reference-style cuSPARSE/cuBLAS baseline written by this project; "
                              + ("faithful: per-iteration workspace malloc" if faithful else "hoisted workspace") + ")",
                      "r0norm": float(np.linalg.norm(prob.rhs("random", seed=1))) if n <= 1 << 24 else None,
                      "rnorm_after_steps": rn, "rnorm_first": hist[:3].tolist(),
                      "config": {"model": f"{a.problem} {'x'.join(map(str, grid))} fp64 unpreconditioned CG",
                                 "nnz": nnz, "clocks_note": "sampled over the whole baseline call (assembly included)"}}))
    return 0
