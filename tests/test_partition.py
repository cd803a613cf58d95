"""Tiers T3/T7 on the CPU: partition bookkeeping and partition invariance through the fake fabric."""
import numpy as np
import pytest

from cuda_mpi_parallel_b200 import Problem, cpu_cg, make_layout, make_partition
from cuda_mpi_parallel_b200.parallel.partition import gather_global, scatter_global


@pytest.mark.parametrize("prob,nranks,decomp", [
    (Problem.poisson2d(10, 13), 4, 1), (Problem.poisson3d(6, 9, 11), 4, 1), (Problem.poisson3d(6, 9, 11), 6, 2),
    (Problem.stencil27(5, 8, 8), 8, 2)])
def test_partitions_tile_the_grid(prob, nranks, decomp):
    gx, gy, gz = prob.dims
    seen = np.zeros((gz, gy, gx), dtype=int)
    parts = [make_partition(prob, nranks, r, decomp) for r in range(nranks)]
    for p in parts:
        seen[p.z0:p.z0 + p.nz, p.y0:p.y0 + p.ny, :] += 1
        lay = make_layout(prob, p)
        assert lay.pitch % 16 == 0 and lay.pitch >= p.nx and lay.vec_elems % 32 == 0
        assert lay.owned == p.nx * p.ny * p.nz
    assert (seen == 1).all()
    # neighbour symmetry: if q is my neighbour in direction d, I am q's neighbour in 8-d
    for p in parts:
        for d in range(9):
            q = p.nbr[d]
            if q >= 0:
                assert parts[q].nbr[8 - d] == p.rank


def test_edge_neighbours_only_for_27pt():
    p7 = make_partition(Problem.poisson3d(8, 8, 8), 4, 0, 2)
    p27 = make_partition(Problem.stencil27(8, 8, 8), 4, 0, 2)
    diag = [0, 2, 6, 8]
    assert all(p7.nbr[d] < 0 for d in diag)
    assert any(p27.nbr[d] >= 0 for d in diag)


def test_scatter_gather_roundtrip():
    prob = Problem.poisson3d(4, 6, 7)
    g = np.arange(prob.n, dtype=float)
    parts = [make_partition(prob, 6, r, 2) for r in range(6)]
    locs = [scatter_global(prob, p, g) for p in parts]
    np.testing.assert_array_equal(gather_global(prob, parts, locs), g)


@pytest.mark.parametrize("prob,configs", [
    (Problem.poisson2d(64, 64), [(2, 1), (4, 1), (8, 1)]),                   # BASELINE config 1 (2 ranks)
    (Problem.poisson3d(12, 12, 16), [(2, 1), (8, 1), (4, 2), (8, 2)]),
    (Problem.stencil27(10, 12, 12), [(2, 1), (4, 2), (6, 2), (8, 2)]),
])
@pytest.mark.parametrize("variant", ["classic", "pipelined"])
def test_partition_invariance(prob, configs, variant):
    b = prob.rhs("random", seed=11)
    ref = cpu_cg(prob, b, tol=1e-8, variant=variant)
    assert ref.converged
    for nranks, decomp in configs:
        s = cpu_cg(prob, b, tol=1e-8, nranks=nranks, decomp=decomp, variant=variant)
        assert abs(s.iters - ref.iters) <= 1
        n = min(s.iters, ref.iters)
        big = ref.history[:n] > 1e-6 * ref.r0norm
        np.testing.assert_allclose(s.history[:n][big], ref.history[:n][big], rtol=1e-8)
        np.testing.assert_allclose(s.x, ref.x, atol=1e-7)


def test_csr_row_block_partition():
    p = Problem.poisson2d(7, 5)
    rowptr, colind, val = p.to_csr()
    pc = Problem.csr(rowptr, colind, val)
    parts = [make_partition(pc, 4, r) for r in range(4)]
    assert [q.x0 for q in parts] == [0, 9, 18, 27] and [q.nx for q in parts] == [9, 9, 9, 8]
    g = np.arange(pc.n, dtype=float)
    np.testing.assert_array_equal(gather_global(pc, parts, [scatter_global(pc, q, g) for q in parts]), g)
