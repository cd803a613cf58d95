from .problems import Problem
from .cpu import cpu_cg, cpu_spmv
from .solver import CGSolver, SolveResult

__all__ = ["Problem", "cpu_cg", "cpu_spmv", "CGSolver", "SolveResult"]
