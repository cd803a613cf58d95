"""CPU oracle (test tiers T1/T7): the reference algorithm restated in plain C++ fp64.

``cpu_cg`` follows CUDACG.cu:248-352 exactly (x0 = 0, r = p = b, absolute ``||r||_2 < tol`` after the
x/r update).  For stencil problems it can run ``nranks`` emulated ranks in one process so that the
partition / halo / rank-ordered reduction logic is tested without a GPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from .. import _native as N
from .problems import Problem


@dataclass
class CpuSolve:
    x: np.ndarray
    iters: int
    converged: bool
    r0norm: float
    rnorm: float
    history: np.ndarray


def cpu_cg(problem: Problem, b: np.ndarray, tol: float = 1e-7, maxit: int = 2000, tol_mode: str = "abs",
           nranks: int = 1, decomp: int = 1, Py: int = 0, Pz: int = 0, variant: str = "classic") -> CpuSolve:
    b = np.ascontiguousarray(b, dtype=np.float64)
    if b.size != problem.n:
        raise ValueError(f"b has {b.size} entries, problem has {problem.n}")
    x = np.zeros(problem.n)
    hist = np.zeros(maxit + 1)
    res = N.CpuResult()
    tm = N.TOL_REL if tol_mode == "rel" else N.TOL_ABS
    L = N.lib()
    if problem.is_csr:
        if variant != "classic" or nranks != 1:
            raise ValueError("CSR oracle is classic CG on one rank")
        rowptr = problem.rowptr.astype(np.int64)
        N.check(L.cgb_cpu_cg_csr(problem.n, rowptr.ctypes.data, problem.colind.ctypes.data, problem.val.ctypes.data,
                                 b.ctypes.data, x.ctypes.data, tol, tm, maxit, hist.ctypes.data, C.byref(res)))
    else:
        v = {"classic": 0, "pipelined": 1}[variant]
        N.check(L.cgb_cpu_cg_stencil(C.byref(problem.stencil), nranks, decomp, Py, Pz, b.ctypes.data, x.ctypes.data,
                                     tol, tm, maxit, v, hist.ctypes.data, C.byref(res)))
    return CpuSolve(x, res.iters, bool(res.converged), res.r0norm, res.rnorm, hist[:res.iters].copy())


def cpu_spmv(problem: Problem, x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float64)
    if problem.is_csr:
        return problem.to_scipy() @ x
    y = np.zeros(problem.n)
    N.check(N.lib().cgb_cpu_spmv_stencil(C.byref(problem.stencil), x.ctypes.data, y.ctypes.data))
    return y
