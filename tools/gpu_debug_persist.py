"""Debug aid: one tiny persistent-kernel solve with short device timeouts; prints the device state afterwards."""
import os, sys
os.environ.setdefault("CGB_TIMEOUT_MS", "200")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from cuda_mpi_parallel_b200 import CGSolver, Problem, cpu_cg, _native as N

dims = [int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "64x48x40").split("x")]
prob = Problem.poisson3d(*dims) if len(dims) == 3 else Problem.poisson2d(*dims)
b = prob.rhs("random", seed=7)
s = CGSolver(prob)
s.set_rhs(b)
r = s.solve(fixed_iters=5, chunk=5, variant="merged", persistent=False)
print("two-kernel merged:", r, flush=True)
N.lib().cgb_ctx_debug_print(s._ctx)
for fixed, chunk in ((2, 2), (5, 3), (12, 5)):
    try:
        r = s.solve(fixed_iters=fixed, chunk=chunk, variant="merged", persistent=True)
        print("fixed", fixed, "chunk", chunk, "->", r, flush=True)
    except Exception as e:
        print("EXC", e, flush=True)
        N.lib().cgb_ctx_debug_print(s._ctx)
        sys.exit(1)
    N.lib().cgb_ctx_debug_print(s._ctx)
ref = cpu_cg(prob, b, tol=1e-8)
r = s.solve(tol=1e-8, variant="merged", persistent=True, chunk=7)
print("converge:", r, "ref iters", ref.iters, "x err", float(np.abs(s.x().cpu().numpy() - ref.x).max()), flush=True)
N.lib().cgb_ctx_debug_print(s._ctx)
