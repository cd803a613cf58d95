"""Multi-process CPU path (BASELINE config 1: "2D 5-point Poisson 64x64, 2 ranks, CPU reference path").

One process per rank over a ``torch.distributed`` group (``gloo`` on machines without a GPU).  Each rank
owns the same padded box the GPU path uses; halos travel as point-to-point messages and scalars are
all-gathered and summed in rank order, so every rank holds bit-identical alpha/beta — the property the
GPU fabric's one-shot all-reduce provides on the device.  Algorithm and stopping rule restate
CUDACG.cu:248-352.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from .. import _native as N
from .partition import make_layout, make_partition


@dataclass
class DistSolve:
    x_local: np.ndarray
    iters: int
    converged: bool
    r0norm: float
    rnorm: float
    history: np.ndarray


def _dir(d):
    return d % 3 - 1, d // 3 - 1


class _Box:
    def __init__(self, problem, part, lay):
        self.problem, self.part, self.lay = problem, part, lay
        self.shape = (lay.pz, lay.py, lay.pitch)

    def zeros(self):
        return np.zeros(int(self.lay.vec_elems))

    def view(self, v):
        L = self.lay
        return v[:L.pitch * L.py * L.pz].reshape(self.shape)

    def owned(self, v):
        L = self.lay
        return self.view(v)[L.gz:L.gz + L.nz, L.gy:L.gy + L.ny, :L.nx]


def _allsum(vals, group):
    """Sum of per-rank scalars in rank order (bit-identical on every rank)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mine = torch.tensor(vals, dtype=torch.float64)
    allv = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine, group=group)
    out = np.zeros(len(vals))
    for t in allv:
        out += t.numpy()
    return out


def _halo_exchange(box: _Box, v, parts, group):
    import torch
    import torch.distributed as dist
    L, part = box.lay, box.part
    V = box.view(v)
    reqs, recvs = [], []
    for d in range(N.NDIR):
        q = part.nbr[d]
        if q < 0:
            continue
        dy, dz = _dir(d)
        ys = slice(L.gy, L.gy + L.ny) if dy == 0 else (slice(L.gy, L.gy + 1) if dy < 0 else slice(L.gy + L.ny - 1, L.gy + L.ny))
        zs = slice(L.gz, L.gz + L.nz) if dz == 0 else (slice(L.gz, L.gz + 1) if dz < 0 else slice(L.gz + L.nz - 1, L.gz + L.nz))
        send = torch.from_numpy(np.ascontiguousarray(V[zs, ys, :L.nx]))
        reqs.append(dist.isend(send, q, group=group, tag=d))
        # what I receive from q was sent in q's direction 8-d; it lands in my ghost layer on side d
        gy = slice(L.gy, L.gy + L.ny) if dy == 0 else (slice(0, 1) if dy < 0 else slice(L.gy + L.ny, L.gy + L.ny + 1))
        gz = slice(L.gz, L.gz + L.nz) if dz == 0 else (slice(0, 1) if dz < 0 else slice(L.gz + L.nz, L.gz + L.nz + 1))
        buf = torch.empty(V[gz, gy, :L.nx].shape, dtype=torch.float64)
        reqs.append(dist.irecv(buf, q, group=group, tag=8 - d))
        recvs.append((gz, gy, buf))
    for r in reqs:
        r.wait()
    for gz, gy, buf in recvs:
        V[gz, gy, :L.nx] = buf.numpy()


def _cpu_dist_cg_csr(problem, b_local, group, tol, maxit, tol_mode) -> DistSolve:
    """General CSR (the operand of CUDACG.cu:213-216) in row blocks: there is no geometric halo, so every iteration
    all-gathers p (rank order = row order) and multiplies the rank's rows by it - the CPU counterpart of the GPU
    path's remote loads of p (cgb_ctx_create_csr_dist)."""
    import scipy.sparse as sp
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    part = make_partition(problem, world, rank, 1)
    r0_, r1_ = part.x0, part.x0 + part.nx
    e0, e1 = int(problem.rowptr[r0_]), int(problem.rowptr[r1_])
    A = sp.csr_matrix((problem.val[e0:e1], problem.colind[e0:e1], problem.rowptr[r0_:r1_ + 1] - e0),
                      shape=(part.nx, problem.n))
    counts = [make_partition(problem, world, q, 1).nx for q in range(world)]

    cmax = max(counts)

    def gather(v):
        mine = torch.zeros(cmax, dtype=torch.float64)       # row blocks differ by at most one row: pad to the largest
        mine[:part.nx] = torch.from_numpy(np.ascontiguousarray(v))
        parts = [torch.empty(cmax, dtype=torch.float64) for _ in counts]
        dist.all_gather(parts, mine, group=group)
        return np.concatenate([t.numpy()[:c] for t, c in zip(parts, counts)])

    x = np.zeros(part.nx)
    r = np.array(b_local, dtype=np.float64).reshape(-1)
    if r.size != part.nx:
        raise ValueError(f"rank {rank} owns {part.nx} rows, b_local has {r.size}")
    p = r.copy()
    rn = float(np.sqrt(_allsum([float(r @ r)], group)[0]))
    rho, r0 = rn * rn, rn
    thresh = tol * rn if tol_mode == "rel" else tol
    hist, iters, conv = [], 0, False
    for it in range(maxit):
        Ap = A @ gather(p)
        alpha = rho / _allsum([float(p @ Ap)], group)[0]
        x += alpha * p
        r -= alpha * Ap
        rhop = rho
        rn = float(np.sqrt(_allsum([float(r @ r)], group)[0]))
        hist.append(rn)
        iters = it + 1
        if rn < thresh:
            conv = True
            break
        rho = rn * rn
        p = r + (rho / rhop) * p
    return DistSolve(x, iters, conv, r0, rn, np.array(hist))


def cpu_dist_cg(problem, b_local: np.ndarray, group=None, tol: float = 1e-7, maxit: int = 2000,
                tol_mode: str = "abs", decomp: int = 1) -> DistSolve:
    import torch.distributed as dist
    if problem.is_csr:
        return _cpu_dist_cg_csr(problem, b_local, group, tol, maxit, tol_mode)
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    part = make_partition(problem, world, rank, decomp)
    lay = make_layout(problem, part)
    box = _Box(problem, part, lay)
    Lc = N.lib()
    x, r, p, Ap = box.zeros(), box.zeros(), box.zeros(), box.zeros()
    box.owned(r)[...] = np.asarray(b_local, dtype=np.float64).reshape(lay.nz, lay.ny, lay.nx)
    box.owned(p)[...] = box.owned(r)
    rn = float(np.sqrt(_allsum([float((box.owned(r) ** 2).sum())], group)[0]))
    rho = rn * rn
    r0 = rn
    thresh = tol * rn if tol_mode == "rel" else tol
    hist, iters, conv = [], 0, False
    d = C.c_double()
    for it in range(maxit):
        _halo_exchange(box, p, None, group)
        N.check(Lc.cgb_cpu_spmv_local(C.byref(problem.stencil), C.byref(lay), p.ctypes.data, Ap.ctypes.data, C.byref(d)))
        pAp = _allsum([d.value], group)[0]
        alpha = rho / pAp
        box.owned(x)[...] += alpha * box.owned(p)
        box.owned(r)[...] -= alpha * box.owned(Ap)
        rhop = rho
        rn = float(np.sqrt(_allsum([float((box.owned(r) ** 2).sum())], group)[0]))
        hist.append(rn)
        iters = it + 1
        if rn < thresh:
            conv = True
            break
        rho = rn * rn
        beta = rho / rhop
        box.owned(p)[...] = box.owned(r) + beta * box.owned(p)
    return DistSolve(np.ascontiguousarray(box.owned(x)).reshape(-1), iters, conv, r0, rn, np.array(hist))
