"""GPU solver: the user-facing object around one per-rank native context.

What the reference does inline in ``main`` — stage A and b (CUDACG.cu:120-153), run the loop
(:244-352), read x back (:355) — is the ``CGSolver`` life cycle::

    s = CGSolver(Problem.poisson3d(512))            # one rank, current CUDA device
    s.set_rhs("ones")                               # or a numpy / torch array (host or device)
    res = s.solve(tol=1e-7, maxit=2000)             # device-resident loop, graphs, no host sync
    x = s.x()                                       # torch tensor on the device

Multi-GPU: build one solver per rank (``rank``/``nranks``), then
:func:`cuda_mpi_parallel_b200.parallel.attach_torch_distributed` (torchrun) or
``attach_single_process``.  Scalars are bit-identical on all ranks, so every rank returns the same
iteration count and residual history.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np

from .. import _native as N
from ..parallel.partition import make_layout, make_partition
from .problems import Problem, _RHS

_VARIANTS = {"auto": N.VAR_AUTO, "fused": N.VAR_FUSED, "recompute": N.VAR_MERGED, "merged": N.VAR_MERGED2, "pipelined": N.VAR_PIPELINED}
_SPMV = {"auto": N.SPMV_AUTO, "simple": N.SPMV_SIMPLE, "tma": N.SPMV_TMA}


@dataclass
class SolveResult:
    iters: int
    status: int            # 1 converged, 2 maxit, 0 running, <0 error
    r0norm: float
    rnorm: float
    device_ms: float
    launches: int

    @property
    def converged(self) -> bool:
        return self.status == 1

    @property
    def iters_per_s(self) -> float:
        return self.iters / (self.device_ms * 1e-3) if self.device_ms > 0 else float("nan")


def _ptr(t) -> int:
    """Device/host pointer of a torch tensor or numpy array."""
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


class CGSolver:
    def __init__(self, problem: Problem, rank: int = 0, nranks: int = 1, decomp: int = 1, Py: int = 0,
                 Pz: int = 0, device: Optional[int] = None, maxit_cap: int = 10000):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("CGSolver needs a CUDA device (use cpu_cg for the CPU oracle)")
        self.problem = problem
        self.rank, self.nranks = rank, nranks
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.part = make_partition(problem, nranks, rank, decomp, Py, Pz)
        self.layout = make_layout(problem, self.part)
        self.maxit_cap = maxit_cap
        L = N.lib()
        ctx = C.c_void_p()
        if problem.is_csr and nranks > 1:
            # row-block distributed CSR: this rank's rows, local row offsets, global column indices
            r0, r1 = self.part.x0, self.part.x0 + self.part.nx
            e0, e1 = int(problem.rowptr[r0]), int(problem.rowptr[r1])
            self._csr_local = (np.ascontiguousarray(problem.rowptr[r0:r1 + 1] - e0, dtype=np.int32),
                               np.ascontiguousarray(problem.colind[e0:e1]), np.ascontiguousarray(problem.val[e0:e1]))
            rp, ci, va = self._csr_local
            N.check(L.cgb_ctx_create_csr_dist(problem.n, C.byref(self.part), e1 - e0, rp.ctypes.data, ci.ctypes.data,
                                              va.ctypes.data, self.device, maxit_cap, C.byref(ctx)))
        elif problem.is_csr:
            N.check(L.cgb_ctx_create_csr(problem.n, int(problem.val.size), problem.rowptr.ctypes.data,
                                         problem.colind.ctypes.data, problem.val.ctypes.data, self.device,
                                         maxit_cap, C.byref(ctx)))
        else:
            N.check(L.cgb_ctx_create(C.byref(problem.stencil), C.byref(self.part), self.device, maxit_cap,
                                     C.byref(ctx)))
        self._ctx = ctx
        self._opts = N.SolveOpts()
        L.cgb_solve_opts_default(C.byref(self._opts))
        if nranks == 1:
            N.check(L.cgb_ctx_fabric_reset(self._ctx))

    # ---- life cycle ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None):
            N.lib().cgb_ctx_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_local(self) -> int:
        return int(self.layout.owned)

    def _stream(self, stream=None) -> Optional[int]:
        import torch
        if stream is None:
            stream = torch.cuda.current_stream(self.device)
        if stream.cuda_stream == 0:
            # The legacy default stream cannot be captured into graphs: the context then runs on its
            # private stream, made consistent with torch's work by a synchronise on both sides.
            stream.synchronize()
        return C.c_void_p(stream.cuda_stream)

    # ---- data ---------------------------------------------------------------------------
    def set_rhs(self, b="ones", seed: int = 0, stream=None) -> None:
        """b: generator name ('ones'|'random'|'manufactured'), or this rank's owned box as a numpy
        array / torch tensor (host pinned or device), compact with x fastest."""
        L = N.lib()
        s = self._stream(stream)
        if isinstance(b, str):
            N.check(L.cgb_ctx_set_rhs_generated(self._ctx, _RHS[b], seed, s))
            return
        if isinstance(b, np.ndarray):
            b = np.ascontiguousarray(b, dtype=np.float64)
            if b.size != self.n_local:
                raise ValueError(f"rhs has {b.size} entries, rank owns {self.n_local}")
            N.check(L.cgb_ctx_set_rhs(self._ctx, b.ctypes.data, 0, s))
            import torch
            torch.cuda.current_stream(self.device).synchronize()  # pageable source must outlive the copy
            return
        import torch
        assert b.dtype == torch.float64 and b.is_contiguous() and b.numel() == self.n_local
        N.check(L.cgb_ctx_set_rhs(self._ctx, C.c_void_p(b.data_ptr()), 1 if b.is_cuda else 0, s))
        self._keep = b  # keep the source alive until the stream has consumed it

    def set_x0(self, x0, stream=None) -> None:
        """Initial guess for ``solve(warm_start=True)``: this rank's owned box (numpy or torch, host or device)."""
        L = N.lib()
        s = self._stream(stream)
        if isinstance(x0, np.ndarray):
            x0 = np.ascontiguousarray(x0, dtype=np.float64)
            assert x0.size == self.n_local
            N.check(L.cgb_ctx_set_x0(self._ctx, x0.ctypes.data, 0, s))
            import torch
            torch.cuda.current_stream(self.device).synchronize()
            return
        import torch
        assert x0.dtype == torch.float64 and x0.is_contiguous() and x0.numel() == self.n_local
        N.check(L.cgb_ctx_set_x0(self._ctx, C.c_void_p(x0.data_ptr()), 1 if x0.is_cuda else 0, s))
        self._keep_x0 = x0

    # Checkpoint = the iterate (plus bookkeeping). Resuming restarts CG from it (r0 = b - A x): the
    # Krylov space is rebuilt, which is the standard restart of an unpreconditioned CG. The reference
    # has no checkpointing at all (nothing is ever written to disk).
    def save_checkpoint(self, path, iters_done: int = 0) -> None:
        np.savez(path, x=self.x().cpu().numpy(), rank=self.rank, nranks=self.nranks, iters_done=iters_done,
                 dims=np.array(self.problem.dims), box=np.array([self.part.y0, self.part.ny, self.part.z0, self.part.nz]))

    def load_checkpoint(self, path) -> int:
        ck = np.load(path)
        if int(ck["rank"]) != self.rank or int(ck["nranks"]) != self.nranks or tuple(ck["dims"]) != tuple(self.problem.dims):
            raise ValueError("checkpoint does not belong to this rank/partition/problem")
        self.set_x0(ck["x"])
        return int(ck["iters_done"])

    def x(self, out=None, stream=None):
        """Owned part of the solution as a device tensor (compact, x fastest)."""
        import torch
        if out is None:
            out = torch.empty(self.n_local, dtype=torch.float64, device=f"cuda:{self.device}")
        N.check(N.lib().cgb_ctx_get_x(self._ctx, C.c_void_p(out.data_ptr()), 1 if out.is_cuda else 0,
                                      self._stream(stream)))
        return out

    def set_host_history(self, pinned=None) -> None:
        """Stream ||r||_2 of every iteration into a pinned host tensor (device-side zero-copy store, no host
        stall, no extra stream op): the reference reads this scalar back every iteration (CUDACG.cu:328)
        but blocks the host to do it. Pass None to switch off. Call before solve()/begin()."""
        if pinned is None:
            N.check(N.lib().cgb_ctx_set_host_history(self._ctx, None, 0))
            self._host_hist = None
            return
        assert pinned.is_pinned() and pinned.dtype.is_floating_point and pinned.element_size() == 8
        N.check(N.lib().cgb_ctx_set_host_history(self._ctx, C.c_void_p(pinned.data_ptr()), pinned.numel()))
        self._host_hist = pinned

    def set_trace(self, buf=None) -> None:
        """Per-iteration device timestamps of the persistent kernel (profiling aid): ``buf`` is a CUDA int64 tensor
        with 4 entries per iteration (ns of %globaltimer: p.Ap barrier reached / alpha released / r.r barrier
        reached / beta released), or None to switch off. Call before begin()/solve()."""
        if buf is None:
            N.check(N.lib().cgb_ctx_set_trace(self._ctx, None, 0))
            self._trace = None
            return
        assert buf.is_cuda and buf.element_size() == 8 and buf.is_contiguous()
        N.check(N.lib().cgb_ctx_set_trace(self._ctx, C.c_void_p(buf.data_ptr()), buf.numel() // 4))
        self._trace = buf

    def history(self, n: int) -> np.ndarray:
        h = np.zeros(max(n, 0))
        N.check(N.lib().cgb_ctx_get_history(self._ctx, h.ctypes.data, int(n)))
        return h

    # ---- solve --------------------------------------------------------------------------
    def _fill_opts(self, tol, maxit, tol_mode, variant, spmv, use_graph, chunk, fixed_iters, stream, warm_start=False,
                   persistent=None):
        o = self._opts
        o.tol = tol
        o.maxit = maxit
        o.tol_mode = N.TOL_REL if tol_mode == "rel" else N.TOL_ABS
        o.variant = _VARIANTS[variant]
        o.spmv_impl = _SPMV[spmv]
        o.use_graph = 2 if use_graph == "while" else int(bool(use_graph))
        o.chunk = chunk
        o.fixed_iters = fixed_iters
        o.warm_start = int(bool(warm_start))
        o.persistent = 0 if persistent is None else (1 if persistent else -1)
        o.stream = self._stream(stream)
        return o

    def solve(self, tol: float = 1e-7, maxit: int = 2000, tol_mode: str = "abs", variant: str = "auto",
              spmv: str = "auto", use_graph: bool = True, chunk: int = 25, fixed_iters: int = 0,
              stream=None, warm_start: bool = False, persistent=None) -> SolveResult:
        """warm_start: start from the x the solver holds (``set_x0`` / ``load_checkpoint`` / the previous solve)
        instead of the reference's x0 = 0: r0 = b - A x0.
        persistent (merged variant): None = two kernels per iteration unless CGB_PERSISTENT=1 is set,
        True = require the persistent cooperative kernel (one launch per chunk), False = never use it."""
        o = self._fill_opts(tol, maxit, tol_mode, variant, spmv, use_graph, chunk, fixed_iters, stream, warm_start,
                            persistent)
        r = N.SolveResult()
        N.check(N.lib().cgb_solve(self._ctx, C.byref(o), C.byref(r)))
        return SolveResult(r.iters, r.status, r.r0norm, r.rnorm, r.device_ms, r.launches)

    # Split-phase API (bench.py times `iterate` between CUDA events on the caller's stream).
    def begin(self, tol: float = 1e-7, maxit: int = 2000, tol_mode: str = "abs", variant: str = "auto",
              spmv: str = "auto", use_graph: bool = True, chunk: int = 25, fixed_iters: int = 0, stream=None,
              warm_start: bool = False, persistent=None):
        o = self._fill_opts(tol, maxit, tol_mode, variant, spmv, use_graph, chunk, fixed_iters, stream, warm_start,
                            persistent)
        N.check(N.lib().cgb_solve_begin(self._ctx, C.byref(o)))

    def prepare(self, **kw):
        """Allocate and capture everything `begin(**kw)` would, without launching. Needed only when one
        host thread drives several ranks (see cgb_solve_prepare)."""
        args = dict(tol=1e-7, maxit=2000, tol_mode="abs", variant="auto", spmv="auto", use_graph=True, chunk=25,
                    fixed_iters=0, stream=None, warm_start=False, persistent=None)
        args.update(kw)
        o = self._fill_opts(**args)
        N.check(N.lib().cgb_solve_prepare(self._ctx, C.byref(o)))

    def iterate(self, n: int):
        N.check(N.lib().cgb_solve_iterate(self._ctx, int(n)))

    def iterate_streamed(self, n: int, host_hist):
        """n iterations; after each one ||r||_2 is copied (stream-ordered, 8 bytes) into the pinned
        host tensor ``host_hist[i]`` — the reference reads this scalar back every iteration
        (CUDACG.cu:328) but blocks the host to do so."""
        assert host_hist.is_pinned() and host_hist.numel() >= n
        N.check(N.lib().cgb_solve_iterate_streamed(self._ctx, int(n), C.c_void_p(host_hist.data_ptr())))

    def poll(self) -> SolveResult:
        r = N.SolveResult()
        N.check(N.lib().cgb_solve_poll(self._ctx, 1, C.byref(r)))
        return SolveResult(r.iters, r.status, r.r0norm, r.rnorm, r.device_ms, r.launches)

    def finish(self) -> SolveResult:
        r = N.SolveResult()
        N.check(N.lib().cgb_solve_finish(self._ctx, C.byref(r)))
        return SolveResult(r.iters, r.status, r.r0norm, r.rnorm, r.device_ms, r.launches)

    @property
    def launch_count(self) -> int:
        return int(N.lib().cgb_ctx_launch_count(self._ctx))
