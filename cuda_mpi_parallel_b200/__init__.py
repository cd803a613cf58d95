"""cuda_mpi_parallel_b200 — B200-native distributed conjugate-gradient framework.

Capabilities of the reference ``CUDACG.cu`` (fp64 unpreconditioned CG on a CSR system, absolute
``||r||_2 < tol`` stopping test, ``%f`` solution printout) re-designed for sm_100a and extended, per
the project's north star, to matrix-free structured Poisson/stencil problems with 1D/2D row-block
domain decomposition over NVLink peer memory.

Layout::

    models/    problems (ref 3x3, Poisson 2D/3D, 27-pt, CSR), CPU oracle, GPU solver classes
    ops/       the hot kernels as standalone ops on torch tensors
    parallel/  partitions, padded layouts, peer-memory fabric bootstrap (torchrun / single process)
    utils/     in-tree sm_100a build (build.py), NVML clock sampling for benchmarks (clocks.py)
"""
from . import _native
from .models.problems import Problem
from .models.cpu import cpu_cg, cpu_spmv
from .models.solver import CGSolver, SolveResult
from .parallel.partition import make_partition, make_layout

__all__ = ["Problem", "cpu_cg", "cpu_spmv", "CGSolver", "SolveResult", "make_partition", "make_layout", "_native"]
__version__ = "0.1.0"
