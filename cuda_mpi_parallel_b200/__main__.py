"""Multi-process command line: one rank per process, launched with torchrun (or plain python for 1 rank).

    python -m cuda_mpi_parallel_b200 --problem poisson3d7 --n 256 --hist-out hist.txt
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m cuda_mpi_parallel_b200 \
        --problem poisson3d7 --grid 512 --variant merged --json-out run.json
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 -m cuda_mpi_parallel_b200 \
        --problem poisson2d5 --grid 64 64 --device cpu       # BASELINE config 1 (gloo, no GPU)
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 -m cuda_mpi_parallel_b200 \
        --problem csr:matrix.mtx --x-out x.txt               # general CSR in row blocks over the GPUs
                                                             # (--device cpu: the same over gloo)

Same controls and output contract as ``bin/cgb200`` (which drives all GPUs from one process): with no
arguments it solves the reference's 3x3 system and prints what CUDACG.cu prints (CUDACG.cu:361-365);
``--hist-out`` gets ``iter ||r||`` lines (``%d %.17e``), ``--json-out`` a one-line summary.  The process
group is used to exchange the 64-byte IPC handles (GPU) or to carry halos and scalars (CPU); the GPU
iteration itself never calls NCCL.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np


def parse(argv=None):
    ap = argparse.ArgumentParser(prog="python -m cuda_mpi_parallel_b200")
    ap.add_argument("--problem", default="ref3x3",
                    help="ref3x3 | poisson2d5 | poisson3d7 | stencil3d27 | csr:<file.mtx> (general CSR from a Matrix "
                         "Market file) | csr-<stencil> (the stencil operator assembled as a general CSR matrix)")
    ap.add_argument("--n", "--grid", dest="n", type=int, nargs="+", default=[64],
                    help="grid extents; use the spelling --grid under torchrun (its parser claims --n*)")
    ap.add_argument("--device", default="cuda", choices=["cuda", "cpu"])
    ap.add_argument("--decomp", type=int, default=1, choices=[1, 2])
    ap.add_argument("--variant", default="auto", choices=["auto", "fused", "recompute", "merged", "pipelined"])
    ap.add_argument("--spmv", default="auto", choices=["auto", "simple", "tma"])
    ap.add_argument("--tol", type=float, default=1e-7)          # CUDACG.cu:245
    ap.add_argument("--tol-mode", default="abs", choices=["abs", "rel"])   # CUDACG.cu:333 is absolute
    ap.add_argument("--maxit", type=int, default=2000)          # CUDACG.cu:244
    ap.add_argument("--rhs", default="ones", help="ones | random[:seed] | manufactured")
    ap.add_argument("--graph", default="chunk", choices=["chunk", "while", "off"])
    ap.add_argument("--chunk", type=int, default=25)
    ap.add_argument("--print-x", action="store_true")
    ap.add_argument("--x-out")
    ap.add_argument("--hist-out")
    ap.add_argument("--json-out")
    return ap.parse_args(argv)


def main(argv=None) -> int:
    a = parse(argv)
    from . import CGSolver, Problem, cpu_cg
    from .parallel.partition import gather_global, make_partition, scatter_global

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    dims = a.n * (3 if len(a.n) == 1 else 1)
    if a.problem.startswith("csr:"):                         # the reference's operand: a general CSR matrix
        prob = Problem.from_matrix_market(a.problem[4:])
    elif a.problem.startswith("csr-"):
        prob = Problem.csr(*Problem.from_name(a.problem[4:], dims).to_csr())
    elif a.problem in ("ref3x3", "poisson2d5", "poisson3d7", "stencil3d27"):
        prob = Problem.from_name(a.problem, dims)
    else:
        raise SystemExit(f"unknown problem {a.problem!r}")
    rhs, _, seed = a.rhs.partition(":")
    # (stencil problems beyond 4 M unknowns generate their right-hand side on the device instead)
    b = (prob.rhs("default") if a.problem == "ref3x3"
         else prob.rhs(rhs, seed=int(seed or 0)) if prob.n <= 1 << 22 or prob.is_csr else None)
    if prob.is_csr and prob.n < world:
        raise SystemExit("general CSR in row blocks needs at least one row per rank")

    import torch.distributed as dist
    if world > 1:
        import torch
        if a.device == "cuda":
            local = int(os.environ.get("LOCAL_RANK", "0"))
            torch.cuda.set_device(local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            dist.init_process_group("gloo")

    part = make_partition(prob, world, rank, a.decomp)
    if a.device == "cpu":
        if world == 1:
            s = cpu_cg(prob, b, tol=a.tol, maxit=a.maxit, tol_mode=a.tol_mode,
                       variant="pipelined" if a.variant == "pipelined" else "classic")
            iters, status, r0, rn, hist, x_local, ms, launches = (s.iters, 1 if s.converged else 2, s.r0norm, s.rnorm,
                                                                  s.history, s.x, 0.0, 0)
        else:
            from .parallel.cpu_dist import cpu_dist_cg
            s = cpu_dist_cg(prob, scatter_global(prob, part, b), tol=a.tol, maxit=a.maxit, tol_mode=a.tol_mode,
                            decomp=a.decomp)
            iters, status, r0, rn, hist, x_local, ms, launches = (s.iters, 1 if s.converged else 2, s.r0norm, s.rnorm,
                                                                  s.history, s.x_local, 0.0, 0)
    else:
        sol = CGSolver(prob, rank=rank, nranks=world, decomp=a.decomp, maxit_cap=a.maxit + 2)
        if world > 1:
            from .parallel import attach_torch_distributed
            attach_torch_distributed(sol)
        if b is None:
            sol.set_rhs(rhs, seed=int(seed or 0))            # generated on the device
        else:
            sol.set_rhs(scatter_global(prob, part, b))       # a row block for CSR, the owned box for stencils
        res = sol.solve(tol=a.tol, maxit=a.maxit, tol_mode=a.tol_mode, variant=a.variant, spmv=a.spmv, chunk=a.chunk,
                        use_graph={"chunk": True, "while": "while", "off": False}[a.graph])
        iters, status, r0, rn, ms, launches = res.iters, res.status, res.r0norm, res.rnorm, res.device_ms, res.launches
        hist = sol.history(min(iters, a.maxit))
        x_local = sol.x().cpu().numpy()
    if status < 0:
        print(f"solve failed: status {status}")          # one line, exit 1: the reference's failure contract
        return 1

    want_x = a.print_x or a.x_out or a.problem == "ref3x3"
    x = None
    if want_x:
        if world > 1:
            parts = [None] * world
            dist.all_gather_object(parts, x_local.tolist())
            if rank == 0:
                ps = [make_partition(prob, world, q, a.decomp) for q in range(world)]
                x = gather_global(prob, ps, [np.array(v) for v in parts])
        else:
            x = np.asarray(x_local)
    if rank == 0:
        if (a.print_x or a.problem == "ref3x3") and x is not None:
            for v in x:
                print("%f" % v)                              # CUDACG.cu:361-364
        if a.x_out and x is not None:
            np.savetxt(a.x_out, x, fmt="%.17e")
        if a.hist_out:
            with open(a.hist_out, "w") as f:
                for i, h in enumerate(np.asarray(hist)[:iters]):
                    f.write("%d %.17e\n" % (i + 1, h))
        if a.json_out:
            its = iters / (ms * 1e-3) if ms > 0 else 0.0
            with open(a.json_out, "w") as f:
                json.dump({"problem": a.problem, "n": prob.n, "ranks": world, "device": a.device, "variant": a.variant,
                           "iters": iters, "status": status, "r0norm": r0, "rnorm": rn, "device_ms": ms,
                           "iters_per_s": its, "gdof_per_s": its * prob.n / 1e9, "launches": launches}, f)
                f.write("\n")
        print("Success")                                     # CUDACG.cu:365
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
