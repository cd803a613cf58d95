"""Tier T2: every sm_100a kernel against a plain fp64 reference of the same op (CPU oracle / torch)."""
import numpy as np
import pytest

from cuda_mpi_parallel_b200 import Problem, cpu_spmv, make_layout, make_partition

pytestmark = pytest.mark.gpu

PROBLEMS = [
    Problem.poisson2d(33, 17), Problem.poisson2d(1100, 70), Problem.poisson2d(512, 130),
    Problem.poisson3d(7, 5, 3), Problem.poisson3d(66, 19, 37), Problem.poisson3d(128, 64, 48),
    Problem.stencil27(9, 6, 5), Problem.stencil27(70, 21, 35), Problem.stencil27(64, 64, 40),
    Problem.poisson3d(1030, 7, 5), Problem.stencil27(1537, 6, 4),      # rows wider than one 512-cell TMA tile
    Problem.poisson2d(2300, 9),                                        # 2D rows wider than one 1024-cell tile
    Problem.poisson2d(256, 40), Problem.poisson2d(1024, 33),           # 2D rows of whole 256-cell boxes: one TMA op per row
    Problem.poisson2d(2304, 21),                                       # ... 3 tiles, the last one mostly out of bounds
    Problem.poisson3d(32, 1280, 24), Problem.stencil27(32, 1280, 24),  # 80 bands x 3 z chunks = 240 items on 148 CTAs
]


def _setup(prob):
    part = make_partition(prob, 1, 0)
    return part, make_layout(prob, part)


@pytest.mark.parametrize("impl", ["simple", "tma"])
@pytest.mark.parametrize("prob", PROBLEMS, ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
def test_spmv_stencil_matches_oracle(prob, impl):
    import torch
    from cuda_mpi_parallel_b200 import ops
    _, lay = _setup(prob)
    rng = np.random.default_rng(1)
    x = rng.standard_normal(prob.n)
    want = cpu_spmv(prob, x)
    xp = ops.pad(lay, torch.from_numpy(x).cuda())
    Ap, d = ops.spmv_stencil(prob, lay, xp, with_dot=True, impl=impl)
    got = ops.unpad(lay, Ap).cpu().numpy()
    scale = np.abs(want).max()
    assert np.abs(got - want).max() <= 1e-13 * scale
    assert abs(d.item() - x @ want) <= 1e-12 * abs(x @ want)
    # padding and ghost cells of the output stay zero
    assert Ap.abs().sum().item() == pytest.approx(np.abs(got).sum(), rel=1e-12)


def test_spmv_csr_and_dot():
    import torch
    from cuda_mpi_parallel_b200 import ops
    prob = Problem.stencil27(11, 9, 8)
    rowptr, colind, val = prob.to_csr()
    rng = np.random.default_rng(2)
    x = rng.standard_normal(prob.n)
    want = cpu_spmv(prob, x)
    y, d = ops.spmv_csr(torch.from_numpy(rowptr.astype(np.int32)).cuda(), torch.from_numpy(colind).cuda(),
                        torch.from_numpy(val).cuda(), torch.from_numpy(x).cuda())
    assert np.abs(y.cpu().numpy() - want).max() <= 1e-13 * np.abs(want).max()
    assert abs(d.item() - x @ want) <= 1e-12 * abs(x @ want)
    a = torch.from_numpy(rng.standard_normal(1_000_003)).cuda()
    b = torch.from_numpy(rng.standard_normal(1_000_003)).cuda()
    ref = (a.cpu().numpy() * b.cpu().numpy()).sum()
    got1, got2 = ops.dot(a, b).item(), ops.dot(a, b).item()
    assert got1 == got2                                  # run-to-run bitwise reproducible
    assert abs(got1 - ref) <= 1e-12 * np.abs(a.cpu().numpy() * b.cpu().numpy()).sum()


@pytest.mark.parametrize("prob", [Problem.poisson2d(33, 17), Problem.poisson3d(66, 19, 37), Problem.stencil27(64, 64, 40)],
                         ids=lambda p: p.name)
def test_fused_update_and_p_update(prob):
    import torch
    from cuda_mpi_parallel_b200 import ops
    _, lay = _setup(prob)
    rng = np.random.default_rng(3)
    vs = {k: rng.standard_normal(prob.n) for k in "pAxr"}
    t = {k: ops.pad(lay, torch.from_numpy(v).cuda()) for k, v in vs.items()}
    alpha, beta = 0.37, -1.25
    rr = ops.fused_update(lay, alpha, t["p"], t["A"], t["x"], t["r"])
    x_want = vs["x"] + alpha * vs["p"]
    r_want = vs["r"] - alpha * vs["A"]
    np.testing.assert_allclose(ops.unpad(lay, t["x"]).cpu().numpy(), x_want, rtol=0, atol=1e-14)
    np.testing.assert_allclose(ops.unpad(lay, t["r"]).cpu().numpy(), r_want, rtol=0, atol=1e-14)
    assert abs(rr.item() - r_want @ r_want) <= 1e-12 * (r_want @ r_want)
    ops.p_update(lay, beta, t["r"], t["p"])
    np.testing.assert_allclose(ops.unpad(lay, t["p"]).cpu().numpy(), r_want + beta * vs["p"], rtol=0, atol=1e-14)
