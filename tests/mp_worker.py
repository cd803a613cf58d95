"""Worker for the multi-process tests (launched with torch.distributed.run).

    mode cpu : gloo, CPU distributed CG (BASELINE config 1) vs the single-process oracle
    mode gpu : nccl bootstrap + cudaIpc fabric, GPU solver vs the oracle (one rank per GPU)
"""
import json
import os
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def main():
    mode, name = sys.argv[1], sys.argv[2]
    dims = [int(v) for v in sys.argv[3].split("x")]
    decomp = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    variant = sys.argv[5] if len(sys.argv) > 5 else "fused"
    # "merged+persist": persistent cooperative kernel; "fused+while": one conditional-WHILE graph per solve
    variant, _, opt = variant.partition("+")
    kw = {}
    if opt == "persist":
        kw["persistent"] = True
    elif opt == "while":
        kw["use_graph"] = "while"
    import torch
    import torch.distributed as dist
    from cuda_mpi_parallel_b200 import Problem, cpu_cg
    from cuda_mpi_parallel_b200.parallel.partition import gather_global, make_partition, scatter_global
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    if name.startswith("csr-"):
        # a general CSR problem (here: an assembled stencil operator) distributed by row blocks
        rowptr, colind, val = Problem.from_name(name[4:], dims).to_csr()
        prob = Problem.csr(rowptr, colind, val)
    else:
        prob = Problem.from_name(name, dims)
    b = prob.rhs("random", seed=21)
    part = make_partition(prob, world, rank, decomp)
    b_local = scatter_global(prob, part, b)
    if mode == "soak":
        # T5: many iterations on a tiny box => thousands of flag handshakes and all-reduces; every rank must
        # see bit-identical scalars and nothing may dead-lock (waits are bounded by CGB_TIMEOUT_MS).
        local_rank = int(os.environ["LOCAL_RANK"])
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from cuda_mpi_parallel_b200 import CGSolver
        from cuda_mpi_parallel_b200.parallel import attach_torch_distributed
        sol = CGSolver(prob, rank=rank, nranks=world, decomp=decomp, maxit_cap=4000)
        attach_torch_distributed(sol)
        ok = True
        for v in ("fused", "pipelined", variant):
            sol.set_rhs(b_local)
            res = sol.solve(fixed_iters=1500, variant=v, chunk=50, **(kw if v == variant else {}))
            h = sol.history(200).tolist()       # before round-off takes over
            g = [None] * world
            dist.all_gather_object(g, (res.iters, res.status, h))
            ok = ok and all(x == g[0] for x in g) and res.iters == 1500 and res.status in (0, 2)
        if rank == 0:
            print("MPRESULT " + json.dumps({"ok": bool(ok), "ranks_identical": bool(ok), "world": world}), flush=True)
        dist.barrier()
        dist.destroy_process_group()
        sys.exit(0 if ok else 1)
    if mode == "cpu":
        dist.init_process_group("gloo")
        from cuda_mpi_parallel_b200.parallel.cpu_dist import cpu_dist_cg
        s = cpu_dist_cg(prob, b_local, tol=1e-8, decomp=decomp)
        iters, hist, x_local, status = s.iters, s.history, s.x_local, int(s.converged)
    else:
        local_rank = int(os.environ["LOCAL_RANK"])
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        from cuda_mpi_parallel_b200 import CGSolver
        from cuda_mpi_parallel_b200.parallel import attach_torch_distributed
        sol = CGSolver(prob, rank=rank, nranks=world, decomp=decomp)
        attach_torch_distributed(sol)
        sol.set_rhs(b_local)
        res = sol.solve(tol=1e-8, chunk=5, variant=variant, **kw)
        # second solve on the same fabric: sequence numbers keep growing, history must repeat bitwise
        h1 = sol.history(res.iters)
        res2 = sol.solve(tol=1e-8, chunk=8, variant=variant, **kw)
        assert res2.iters == res.iters and (sol.history(res.iters) == h1).all(), "repeat solve differs"
        iters, hist, status = res.iters, h1, res.status
        x_local = sol.x().cpu().numpy()
    gathered = [None] * world
    dist.all_gather_object(gathered, (iters, hist.tolist(), x_local.tolist()))
    ok = True
    msg = ""
    if rank == 0:
        ref = cpu_cg(prob, b, tol=1e-8, variant="pipelined" if variant == "pipelined" else "classic")
        htol = 1e-6 if variant == "pipelined" else 1e-8
        parts = [make_partition(prob, world, q, decomp) for q in range(world)]
        x = gather_global(prob, parts, [np.array(g[2]) for g in gathered])
        same = all(g[0] == gathered[0][0] and g[1] == gathered[0][1] for g in gathered)   # bit-identical on all ranks
        n = min(iters, ref.iters)
        big = ref.history[:n] > 1e-6 * ref.r0norm
        herr = float(np.max(np.abs(np.array(hist)[:n][big] / ref.history[:n][big] - 1))) if big.any() else 0.0
        xerr = float(np.abs(x - ref.x).max())
        ok = bool(status == 1 and same and abs(iters - ref.iters) <= 1 and herr < htol and xerr < 1e-6)
        msg = json.dumps({"ok": ok, "iters": iters, "ref_iters": ref.iters, "hist_relerr": herr, "x_err": xerr,
                          "ranks_identical": same, "world": world})
        print("MPRESULT " + msg, flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
