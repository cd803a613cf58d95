"""The hot kernels as standalone ops on torch CUDA tensors (fp64).

Each op calls the same sm_100a kernel the solver uses (through ``cgb_op_*`` of the C ABI) — there is
no PyTorch fallback; on a machine without the native library these raise.

Mapping to the reference's library calls (CUDACG.cu): ``spmv_stencil``/``spmv_csr`` = cusparseSpMV (:288)
+ cublasDdot (:304); ``fused_update`` = 2x cublasDaxpy (:314,:320) + cublasDnrm2 (:328);
``p_update`` = cublasDscal (:342) + cublasDaxpy (:347); ``dot`` = cublasDdot.
"""
from __future__ import annotations

import ctypes as C

from .. import _native as N

_scratch = {}


def _vp(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream(t):
    import torch
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def scratch(device):
    """Per-device scratch (partials + control block) shared by the ops; calls on one device must not
    overlap on different streams."""
    import torch
    key = torch.device(device).index
    if key not in _scratch:
        _scratch[key] = torch.zeros(N.lib().cgb_op_scratch_bytes(), dtype=torch.uint8, device=device)
    return _scratch[key]


def _chk(t, n=None):
    import torch
    assert t.is_cuda and t.dtype == torch.float64 and t.is_contiguous(), "ops need contiguous fp64 CUDA tensors"
    if n is not None:
        assert t.numel() == n, f"expected {n} elements, got {t.numel()}"


def pad(layout: N.Layout, compact):
    import torch
    _chk(compact, layout.owned)
    out = torch.empty(layout.vec_elems, dtype=torch.float64, device=compact.device)
    with torch.cuda.device(compact.device):
        N.check(N.lib().cgb_op_pad(C.byref(layout), _vp(compact), _vp(out), _stream(compact)))
    return out


def unpad(layout: N.Layout, padded):
    import torch
    _chk(padded, layout.vec_elems)
    out = torch.empty(layout.owned, dtype=torch.float64, device=padded.device)
    with torch.cuda.device(padded.device):
        N.check(N.lib().cgb_op_unpad(C.byref(layout), _vp(padded), _vp(out), _stream(padded)))
    return out


def dot(a, b):
    import torch
    _chk(a)
    _chk(b, a.numel())
    out = torch.empty(1, dtype=torch.float64, device=a.device)
    with torch.cuda.device(a.device):
        N.check(N.lib().cgb_op_dot(_vp(a), _vp(b), a.numel(), _vp(out), _vp(scratch(a.device)), _stream(a)))
    return out


def spmv_stencil(problem, layout: N.Layout, p_padded, with_dot=True, impl="auto"):
    """Ap = A p on one padded local box; returns (Ap_padded, p.Ap or None)."""
    import torch
    from ..models.solver import _SPMV
    _chk(p_padded, layout.vec_elems)
    Ap = torch.zeros_like(p_padded)
    d = torch.empty(1, dtype=torch.float64, device=p_padded.device) if with_dot else None
    with torch.cuda.device(p_padded.device):
        N.check(N.lib().cgb_op_spmv_stencil(C.byref(problem.stencil), C.byref(layout), _vp(p_padded), _vp(Ap), _vp(d),
                                            _vp(scratch(p_padded.device)), _SPMV[impl], _stream(p_padded)))
    return Ap, d


def spmv_csr(rowptr, colind, val, x, with_dot=True):
    import torch
    n = rowptr.numel() - 1
    assert rowptr.dtype == torch.int32 and colind.dtype == torch.int32
    _chk(val)
    _chk(x, n)
    y = torch.empty_like(x)
    d = torch.empty(1, dtype=torch.float64, device=x.device) if with_dot else None
    with torch.cuda.device(x.device):
        N.check(N.lib().cgb_op_spmv_csr(n, _vp(rowptr), _vp(colind), _vp(val), _vp(x), _vp(y), _vp(d),
                                        _vp(scratch(x.device)), _stream(x)))
    return y, d


def fused_update(layout: N.Layout, alpha: float, p, Ap, x, r):
    """In place: x += alpha p, r -= alpha Ap; returns r.r (1-element tensor)."""
    import torch
    for t in (p, Ap, x, r):
        _chk(t, layout.vec_elems)
    rr = torch.empty(1, dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        N.check(N.lib().cgb_op_fused_update(C.byref(layout), float(alpha), _vp(p), _vp(Ap), _vp(x), _vp(r), _vp(rr),
                                            _vp(scratch(x.device)), _stream(x)))
    return rr


def p_update(layout: N.Layout, beta: float, r, p):
    """In place: p = r + beta p."""
    import torch
    for t in (r, p):
        _chk(t, layout.vec_elems)
    with torch.cuda.device(p.device):
        N.check(N.lib().cgb_op_p_update(C.byref(layout), float(beta), _vp(r), _vp(p), _vp(scratch(p.device)), _stream(p)))
    return p
