"""ctypes view of ``include/cgb200.h`` (the C ABI of libcgb200.so).

The library is built in-tree for sm_100a by :mod:`cuda_mpi_parallel_b200.utils.build`; if the
shared object is missing or stale it is (re)built on import.  Nothing here falls back to a
PyTorch implementation: on a GPU box a missing library is a hard error.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from .utils import build as _build

MAX_RANKS = 16
NDIR = 9
IPC_HANDLE_BYTES = 64

KIND_CSR, KIND_POISSON2D5, KIND_POISSON3D7, KIND_STENCIL3D27 = 0, 1, 2, 3
RHS_ONES, RHS_RANDOM, RHS_MANUFACTURED = 0, 1, 2
TOL_ABS, TOL_REL = 0, 1
VAR_FUSED, VAR_MERGED, VAR_PIPELINED, VAR_AUTO, VAR_MERGED2 = 0, 1, 2, 3, 4
SPMV_AUTO, SPMV_SIMPLE, SPMV_TMA = 0, 1, 2

STATUS_NAMES = {0: "ok", -1: "invalid argument", -2: "CUDA error", -3: "bad state", -4: "peer wait timeout",
                -5: "breakdown (non-finite scalar)", -6: "unsupported"}


class Stencil(C.Structure):
    _fields_ = [("kind", C.c_int32), ("gx", C.c_int32), ("gy", C.c_int32), ("gz", C.c_int32),
                ("c_center", C.c_double), ("c_face", C.c_double), ("c_edge", C.c_double),
                ("c_corner", C.c_double)]


class Partition(C.Structure):
    _fields_ = [("nranks", C.c_int32), ("rank", C.c_int32), ("Py", C.c_int32), ("Pz", C.c_int32),
                ("ry", C.c_int32), ("rz", C.c_int32), ("nx", C.c_int32), ("y0", C.c_int32),
                ("ny", C.c_int32), ("z0", C.c_int32), ("nz", C.c_int32), ("nbr", C.c_int32 * NDIR), ("x0", C.c_int32)]


class Layout(C.Structure):
    _fields_ = [("nx", C.c_int32), ("ny", C.c_int32), ("nz", C.c_int32), ("gy", C.c_int32),
                ("gz", C.c_int32), ("pitch", C.c_int32), ("py", C.c_int32), ("pz", C.c_int32),
                ("vec_elems", C.c_int64), ("owned", C.c_int64)]

    def index(self, i, j, k):
        return i + self.pitch * ((j + self.gy) + self.py * (k + self.gz))


class CpuResult(C.Structure):
    _fields_ = [("iters", C.c_int32), ("converged", C.c_int32), ("r0norm", C.c_double), ("rnorm", C.c_double)]


class SolveOpts(C.Structure):
    _fields_ = [("tol", C.c_double), ("tol_mode", C.c_int32), ("maxit", C.c_int32), ("variant", C.c_int32),
                ("spmv_impl", C.c_int32), ("use_graph", C.c_int32), ("chunk", C.c_int32),
                ("fixed_iters", C.c_int32), ("warm_start", C.c_int32), ("persistent", C.c_int32), ("stream", C.c_void_p)]


class SolveResult(C.Structure):
    _fields_ = [("iters", C.c_int32), ("status", C.c_int32), ("r0norm", C.c_double), ("rnorm", C.c_double),
                ("device_ms", C.c_double), ("launches", C.c_int64)]


class NativeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"cgb200: {STATUS_NAMES.get(code, code)}: {msg}")
        self.code = code


_lib = None


def _declare(lib):
    P = C.POINTER
    vp, i32, i64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    sig = {
        "cgb_stencil_default": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, P(Stencil)]),
        "cgb_stencil_csr_nnz": (i64, [P(Stencil), i64, i64]),
        "cgb_stencil_to_csr": (C.c_int, [P(Stencil), i64, i64, vp, vp, vp]),
        "cgb_ref3x3": (None, [P(i32), P(i32), vp, vp, vp, vp]),
        "cgb_rhs_value": (dbl, [P(Stencil), C.c_int, C.c_uint64, i64]),
        "cgb_manufactured_u": (dbl, [P(Stencil), i64]),
        "cgb_rhs_fill_box": (C.c_int, [P(Stencil), C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, vp]),
        "cgb_partition_make": (C.c_int, [P(Stencil), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, P(Partition)]),
        "cgb_layout_make": (C.c_int, [P(Stencil), P(Partition), P(Layout)]),
        "cgb_cpu_cg_csr": (C.c_int, [i64, vp, vp, vp, vp, vp, dbl, C.c_int, C.c_int, vp, P(CpuResult)]),
        "cgb_cpu_cg_stencil": (C.c_int, [P(Stencil), C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, dbl, C.c_int,
                                         C.c_int, C.c_int, vp, P(CpuResult)]),
        "cgb_cpu_spmv_local": (C.c_int, [P(Stencil), P(Layout), vp, vp, P(dbl)]),
        "cgb_cpu_spmv_stencil": (C.c_int, [P(Stencil), vp, vp]),
        "cgb_solve_opts_default": (None, [P(SolveOpts)]),
        "cgb_ctx_create": (C.c_int, [P(Stencil), P(Partition), C.c_int, C.c_int, P(vp)]),
        "cgb_ctx_create_csr": (C.c_int, [i32, i64, vp, vp, vp, C.c_int, C.c_int, P(vp)]),
        "cgb_ctx_create_csr_dist": (C.c_int, [i32, P(Partition), i64, vp, vp, vp, C.c_int, C.c_int, P(vp)]),
        "cgb_ctx_destroy": (C.c_int, [vp]),
        "cgb_ctx_ipc_export": (C.c_int, [vp, vp]),
        "cgb_ctx_peer_attach_ipc": (C.c_int, [vp, C.c_int, vp]),
        "cgb_ctx_peer_attach_ptr": (C.c_int, [vp, C.c_int, vp]),
        "cgb_ctx_arena_base": (vp, [vp]),
        "cgb_ctx_arena_bytes": (C.c_size_t, [vp]),
        "cgb_ctx_fabric_reset": (C.c_int, [vp]),
        "cgb_ctx_set_rhs_generated": (C.c_int, [vp, C.c_int, C.c_uint64, vp]),
        "cgb_ctx_set_rhs": (C.c_int, [vp, vp, C.c_int, vp]),
        "cgb_ctx_set_x0": (C.c_int, [vp, vp, C.c_int, vp]),
        "cgb_ctx_get_x": (C.c_int, [vp, vp, C.c_int, vp]),
        "cgb_ctx_get_history": (C.c_int, [vp, vp, C.c_int]),
        "cgb_ctx_set_host_history": (C.c_int, [vp, vp, C.c_int]),
        "cgb_solve": (C.c_int, [vp, P(SolveOpts), P(SolveResult)]),
        "cgb_solve_begin": (C.c_int, [vp, P(SolveOpts)]),
        "cgb_solve_prepare": (C.c_int, [vp, P(SolveOpts)]),
        "cgb_solve_iterate": (C.c_int, [vp, C.c_int]),
        "cgb_solve_iterate_streamed": (C.c_int, [vp, C.c_int, vp]),
        "cgb_solve_poll": (C.c_int, [vp, C.c_int, P(SolveResult)]),
        "cgb_solve_finish": (C.c_int, [vp, P(SolveResult)]),
        "cgb_ctx_set_trace": (C.c_int, [vp, vp, C.c_int]),
        "cgb_ctx_debug_print": (C.c_int, [vp]),
        "cgb_ctx_launch_count": (i64, [vp]),
        "cgb_op_scratch_bytes": (C.c_size_t, []),
        "cgb_op_dot": (C.c_int, [vp, vp, i64, vp, vp, vp]),
        "cgb_op_dot_static": (C.c_int, [vp, vp, i64, vp, vp]),
        "cgb_op_dot_static_max_n": (C.c_longlong, []),
        "cgb_op_spmv_stencil": (C.c_int, [P(Stencil), P(Layout), vp, vp, vp, vp, C.c_int, vp]),
        "cgb_op_spmv_csr": (C.c_int, [i32, vp, vp, vp, vp, vp, vp, vp, vp]),
        "cgb_op_fused_update": (C.c_int, [P(Layout), dbl, vp, vp, vp, vp, vp, vp, vp]),
        "cgb_op_p_update": (C.c_int, [P(Layout), dbl, vp, vp, vp, vp]),
        "cgb_op_pad": (C.c_int, [P(Layout), vp, vp, vp]),
        "cgb_op_unpad": (C.c_int, [P(Layout), vp, vp, vp]),
        "cgb_debug_tma_geometry": (C.c_int, [P(Stencil), P(Layout), C.c_int, C.c_int, C.c_int, C.c_int, P(C.c_int)]),
        "cgb_version": (C.c_char_p, []),
        "cgb_last_error": (C.c_char_p, []),
        "cgb_device_count": (C.c_int, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def lib():
    """The loaded native library (built on first use if needed)."""
    global _lib
    if _lib is None:
        if os.environ.get("CGB200_NO_BUILD") != "1" or not _build.LIB.exists():
            _build.build_core()
        _lib = C.CDLL(str(_build.LIB), mode=C.RTLD_GLOBAL)
        _declare(_lib)
    return _lib


def lib_path() -> Path:
    return _build.LIB


def check(rc: int):
    if rc != 0:
        raise NativeError(rc, lib().cgb_last_error().decode())
    return rc
