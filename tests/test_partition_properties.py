"""Property tests (hypothesis) of the topology layer and of the fake-fabric oracle on random small grids."""
import numpy as np
from hypothesis import given, settings, strategies as st

from cuda_mpi_parallel_b200 import Problem, cpu_cg, make_layout, make_partition

kinds = st.sampled_from(["poisson2d5", "poisson3d7", "stencil3d27"])


@st.composite
def cases(draw):
    kind = draw(kinds)
    nx, ny, nz = draw(st.integers(1, 9)), draw(st.integers(2, 9)), draw(st.integers(2, 9))
    prob = Problem.from_name(kind, [nx, ny, nz])
    gx, gy, gz = prob.dims
    decomp = 1 if kind == "poisson2d5" else draw(st.sampled_from([1, 2]))
    nranks = draw(st.integers(1, 8))
    return prob, nranks, decomp


def _valid(prob, nranks, decomp):
    try:
        make_partition(prob, nranks, 0, decomp)
        return True
    except Exception:
        return False


@given(cases())
@settings(max_examples=80, deadline=None)
def test_partitions_tile_the_grid_and_neighbours_are_symmetric(case):
    prob, nranks, decomp = case
    if not _valid(prob, nranks, decomp):
        return                                # more ranks than rows/planes: rejected, which is fine
    gx, gy, gz = prob.dims
    seen = np.zeros((gz, gy, gx), dtype=int)
    parts = [make_partition(prob, nranks, r, decomp) for r in range(nranks)]
    for p in parts:
        assert p.nx == gx and p.ny >= 1 and p.nz >= 1
        seen[p.z0:p.z0 + p.nz, p.y0:p.y0 + p.ny, :] += 1
        lay = make_layout(prob, p)
        assert lay.owned == p.nx * p.ny * p.nz and lay.py == p.ny + 2 * lay.gy and lay.pz == p.nz + 2 * lay.gz
        for d in range(9):
            q = p.nbr[d]
            if q >= 0:
                assert parts[q].nbr[8 - d] == p.rank
                dy, dz = d % 3 - 1, d // 3 - 1
                assert (parts[q].ry, parts[q].rz) == (p.ry + dy, p.rz + dz)
    assert (seen == 1).all()


@given(cases(), st.integers(0, 2 ** 31))
@settings(max_examples=25, deadline=None)
def test_fake_fabric_is_partition_invariant(case, seed):
    prob, nranks, decomp = case
    if prob.n < 4 or not _valid(prob, nranks, decomp):
        return
    b = np.random.default_rng(seed).uniform(-1, 1, prob.n)
    one = cpu_cg(prob, b, tol=1e-9, maxit=400)
    many = cpu_cg(prob, b, tol=1e-9, maxit=400, nranks=nranks, decomp=decomp)
    assert abs(one.iters - many.iters) <= 1
    np.testing.assert_allclose(many.x, one.x, atol=1e-7 * max(1.0, np.abs(one.x).max()))
