"""Peer-memory fabric bootstrap.

Every rank owns one peer-mappable arena (control block with reduction slots and halo flags, plus the
exchanged vectors).  Bootstrap only tells each rank where its peers' arenas are; afterwards the
kernels talk through plain loads/stores over NVLink and no MPI/NCCL call is on any iteration path.

* :func:`attach_single_process` — all ranks live in this process (one context per GPU); peers are
  reached through ``cudaDeviceEnablePeerAccess`` and raw pointers.
* :func:`attach_torch_distributed` — one process per GPU (torchrun); ``cudaIpc`` handles travel over the
  process group once, at start-up.
"""
from __future__ import annotations

import ctypes as C

from .. import _native as N


def attach_single_process(solvers) -> None:
    L = N.lib()
    for s in solvers:
        for t in solvers:
            if t is s:
                continue
            N.check(L.cgb_ctx_peer_attach_ptr(s._ctx, t.rank, L.cgb_ctx_arena_base(t._ctx)))
    for s in solvers:
        N.check(L.cgb_ctx_fabric_reset(s._ctx))


def attach_torch_distributed(solver, group=None) -> None:
    import torch.distributed as dist
    L = N.lib()
    world = dist.get_world_size(group)
    if world != solver.nranks:
        raise ValueError(f"solver was built for {solver.nranks} ranks, process group has {world}")
    buf = (C.c_ubyte * N.IPC_HANDLE_BYTES)()
    N.check(L.cgb_ctx_ipc_export(solver._ctx, buf))
    handles = [None] * world
    dist.all_gather_object(handles, bytes(buf), group=group)
    for q, h in enumerate(handles):
        if q == solver.rank:
            continue
        hb = (C.c_ubyte * N.IPC_HANDLE_BYTES).from_buffer_copy(h)
        N.check(L.cgb_ctx_peer_attach_ipc(solver._ctx, q, hb))
    N.check(L.cgb_ctx_fabric_reset(solver._ctx))
    dist.barrier(group=group)  # nobody pushes before every control block is zeroed
