"""Problem layer: what system is being solved.

The reference hard-codes one 3x3 CSR system in ``main`` (CUDACG.cu:74-117, b at :137-140).  Here a
problem is a value: that same system (:meth:`Problem.ref3x3`), a general CSR matrix, or one of the
matrix-free stencils named by BASELINE.json (2D 5-point, 3D 7-point, 3D 27-point).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional

import numpy as np

from .. import _native as N

_KINDS = {"csr": N.KIND_CSR, "poisson2d5": N.KIND_POISSON2D5, "poisson3d7": N.KIND_POISSON3D7,
          "stencil3d27": N.KIND_STENCIL3D27}
_RHS = {"ones": N.RHS_ONES, "random": N.RHS_RANDOM, "manufactured": N.RHS_MANUFACTURED}


@dataclass
class Problem:
    name: str
    stencil: N.Stencil
    rowptr: Optional[np.ndarray] = None   # CSR problems only
    colind: Optional[np.ndarray] = None
    val: Optional[np.ndarray] = None
    b_default: Optional[np.ndarray] = field(default=None, repr=False)

    # ---- constructors -------------------------------------------------------------------
    @staticmethod
    def _stencil(kind: int, gx: int, gy: int, gz: int) -> N.Stencil:
        st = N.Stencil()
        N.check(N.lib().cgb_stencil_default(kind, gx, gy, gz, C.byref(st)))
        return st

    @classmethod
    def poisson2d(cls, nx: int, ny: Optional[int] = None) -> "Problem":
        """2D 5-point Poisson on an nx x ny grid. Internally the grid is nx x 1 x ny (one row per plane), so
        ``dims`` reports (nx, 1, ny) and the row-block decomposition is the z-slab decomposition."""
        ny = nx if ny is None else ny
        return cls("poisson2d5", cls._stencil(N.KIND_POISSON2D5, nx, ny, 1))

    @classmethod
    def poisson3d(cls, nx: int, ny: Optional[int] = None, nz: Optional[int] = None) -> "Problem":
        ny = nx if ny is None else ny
        nz = nx if nz is None else nz
        return cls("poisson3d7", cls._stencil(N.KIND_POISSON3D7, nx, ny, nz))

    @classmethod
    def stencil27(cls, nx: int, ny: Optional[int] = None, nz: Optional[int] = None) -> "Problem":
        ny = nx if ny is None else ny
        nz = nx if nz is None else nz
        return cls("stencil3d27", cls._stencil(N.KIND_STENCIL3D27, nx, ny, nz))

    @classmethod
    def from_name(cls, name: str, dims) -> "Problem":
        dims = list(dims)
        if name == "ref3x3":
            return cls.ref3x3()
        if name == "poisson2d5":
            return cls.poisson2d(*dims[:2])
        if name == "poisson3d7":
            return cls.poisson3d(*dims[:3])
        if name == "stencil3d27":
            return cls.stencil27(*dims[:3])
        raise ValueError(f"unknown problem {name!r}")

    @classmethod
    def csr(cls, rowptr, colind, val, name: str = "csr") -> "Problem":
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int32)
        colind = np.ascontiguousarray(colind, dtype=np.int32)
        val = np.ascontiguousarray(val, dtype=np.float64)
        n = rowptr.size - 1
        if n < 1 or rowptr[0] != 0 or rowptr[-1] != colind.size or colind.size != val.size:
            raise ValueError("inconsistent CSR arrays")
        return cls(name, cls._stencil(N.KIND_CSR, n, 1, 1), rowptr, colind, val)

    @classmethod
    def from_scipy(cls, A, name: str = "csr") -> "Problem":
        """General CSR problem from any scipy sparse matrix (converted to CSR with sorted int32 indices)."""
        import scipy.sparse as sp
        A = sp.csr_matrix(A, dtype=np.float64)
        if A.shape[0] != A.shape[1]:
            raise ValueError("matrix must be square")
        A.sort_indices()
        return cls.csr(A.indptr, A.indices, A.data, name=name)

    @classmethod
    def from_matrix_market(cls, path, name: str = "csr") -> "Problem":
        """General CSR problem from a MatrixMarket file (symmetric files are expanded), like `cgb200 --problem csr:F`."""
        import scipy.io
        return cls.from_scipy(scipy.io.mmread(str(path)), name=name)

    @classmethod
    def ref3x3(cls) -> "Problem":
        """The reference's system: A=[[3,0,2],[0,2,0],[2,0,1]], b=(3.5,1.5,2), x=(0.5,0.75,1)."""
        n, nnz = C.c_int32(), C.c_int32()
        val = np.zeros(5)
        rowptr = np.zeros(4, dtype=np.int32)
        colind = np.zeros(5, dtype=np.int32)
        b = np.zeros(3)
        N.lib().cgb_ref3x3(C.byref(n), C.byref(nnz), val.ctypes.data, rowptr.ctypes.data, colind.ctypes.data,
                           b.ctypes.data)
        p = cls.csr(rowptr, colind, val, name="ref3x3")
        p.b_default = b
        return p

    # ---- properties ---------------------------------------------------------------------
    @property
    def kind(self) -> int:
        return self.stencil.kind

    @property
    def is_csr(self) -> bool:
        return self.kind == N.KIND_CSR

    @property
    def dims(self):
        return (self.stencil.gx, self.stencil.gy, self.stencil.gz)

    @property
    def n(self) -> int:
        return self.stencil.gx * self.stencil.gy * self.stencil.gz

    # ---- data ---------------------------------------------------------------------------
    def to_csr(self):
        """(rowptr int64, colind int32, val float64) of the assembled operator."""
        if self.is_csr:
            return self.rowptr.astype(np.int64), self.colind, self.val
        L = N.lib()
        nnz = L.cgb_stencil_csr_nnz(C.byref(self.stencil), 0, self.n)
        rowptr = np.zeros(self.n + 1, dtype=np.int64)
        colind = np.zeros(nnz, dtype=np.int32)
        val = np.zeros(nnz, dtype=np.float64)
        N.check(L.cgb_stencil_to_csr(C.byref(self.stencil), 0, self.n, rowptr.ctypes.data, colind.ctypes.data,
                                     val.ctypes.data))
        return rowptr, colind, val

    def to_scipy(self):
        import scipy.sparse as sp
        rowptr, colind, val = self.to_csr()
        return sp.csr_matrix((val, colind, rowptr), shape=(self.n, self.n))

    def rhs(self, kind: str = "ones", seed: int = 0) -> np.ndarray:
        """Global right-hand side (compact, x fastest) from the named generator."""
        if self.is_csr:
            if self.b_default is not None and kind == "default":
                return self.b_default.copy()
            if kind == "ones":
                return np.ones(self.n)
            rng = np.random.default_rng(seed)
            return rng.uniform(-1, 1, self.n)
        L = N.lib()
        k = _RHS[kind]
        if k == N.RHS_ONES:
            return np.ones(self.n)
        out = np.empty(self.n)
        N.check(L.cgb_rhs_fill_box(C.byref(self.stencil), k, seed, 0, self.stencil.gy, 0, self.stencil.gz, out.ctypes.data))
        return out

    def rhs_local(self, part, kind: str = "ones", seed: int = 0, out=None):
        """One rank's owned box of the generated right-hand side (compact, x fastest), bit-identical to the
        device generator ``CGSolver.set_rhs(kind, seed)``. ``out``: optional float64 buffer (numpy array or a
        CPU torch tensor, e.g. pinned) of the box size; a numpy array is returned otherwise."""
        if self.is_csr:
            return self.rhs(kind, seed)[part.x0:part.x0 + part.nx]
        n_local = part.nx * part.ny * part.nz
        if out is None:
            out = np.empty(n_local)
        ptr = out.ctypes.data if isinstance(out, np.ndarray) else out.data_ptr()
        size = out.size if isinstance(out, np.ndarray) else out.numel()
        assert size == n_local, f"buffer holds {size} values, the rank owns {n_local}"
        N.check(N.lib().cgb_rhs_fill_box(C.byref(self.stencil), _RHS[kind], seed, part.y0, part.ny, part.z0, part.nz, ptr))
        return out

    def manufactured_solution(self) -> np.ndarray:
        L = N.lib()
        st = C.byref(self.stencil)
        return np.array([L.cgb_manufactured_u(st, g) for g in range(self.n)])
