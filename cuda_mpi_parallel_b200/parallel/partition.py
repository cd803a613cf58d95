"""Topology layer: 1D/2D row-block partitions and padded local layouts.

The reference has no decomposition (one CSR on device 0, CUDACG.cu:87,213-216).  A partition assigns
each rank a contiguous box of grid rows (z slabs, or y*z pencils for ``decomp=2``; 2D grids split in
y); a layout is that box plus one ghost layer per split-capable axis.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _native as N


def make_partition(problem, nranks: int, rank: int, decomp: int = 1, Py: int = 0, Pz: int = 0) -> N.Partition:
    part = N.Partition()
    N.check(N.lib().cgb_partition_make(C.byref(problem.stencil), nranks, rank, decomp, Py, Pz, C.byref(part)))
    return part


def make_layout(problem, part: N.Partition) -> N.Layout:
    lay = N.Layout()
    N.check(N.lib().cgb_layout_make(C.byref(problem.stencil), C.byref(part), C.byref(lay)))
    return lay


def local_slices(problem, part: N.Partition):
    """Slices selecting a rank's owned box out of a global array shaped (gz, gy, gx)."""
    return (slice(part.z0, part.z0 + part.nz), slice(part.y0, part.y0 + part.ny), slice(0, part.nx))


def scatter_global(problem, part: N.Partition, g: np.ndarray) -> np.ndarray:
    if problem.is_csr:                       # row block of a general CSR problem
        return np.ascontiguousarray(np.asarray(g)[part.x0:part.x0 + part.nx])
    gx, gy, gz = problem.dims
    return np.ascontiguousarray(g.reshape(gz, gy, gx)[local_slices(problem, part)]).reshape(-1)


def gather_global(problem, parts, locals_) -> np.ndarray:
    if problem.is_csr:
        out = np.zeros(problem.n)
        for part, loc in zip(parts, locals_):
            out[part.x0:part.x0 + part.nx] = np.asarray(loc)
        return out
    gx, gy, gz = problem.dims
    out = np.zeros((gz, gy, gx))
    for part, loc in zip(parts, locals_):
        out[local_slices(problem, part)] = np.asarray(loc).reshape(part.nz, part.ny, part.nx)
    return out.reshape(-1)
