"""Tier T1: the CPU oracle against analytic answers and scipy (no GPU)."""
import numpy as np
import pytest

from cuda_mpi_parallel_b200 import Problem, cpu_cg, cpu_spmv


def test_ref3x3_golden_trace():
    # Golden trace of CUDACG.cu:248-352 on the reference's own system.
    p = Problem.ref3x3()
    s = cpu_cg(p, p.rhs("default"), tol=1e-7, maxit=2000)
    assert s.iters == 3 and s.converged
    np.testing.assert_allclose(s.x, [0.5, 0.75, 1.0], atol=1e-12)
    np.testing.assert_allclose(s.history[:2], [8.073855e-01, 1.673428e-01], rtol=1e-6)
    assert s.history[2] < 1e-7
    assert abs(s.r0norm - np.linalg.norm([3.5, 1.5, 2.0])) < 1e-14


@pytest.mark.parametrize("prob", [Problem.poisson2d(9, 7), Problem.poisson3d(5, 6, 4), Problem.stencil27(4, 5, 6)])
def test_stencil_matches_assembled_csr(prob):
    A = prob.to_scipy()
    assert (abs(A - A.T)).max() == 0
    rng = np.random.default_rng(0)
    x = rng.standard_normal(prob.n)
    np.testing.assert_allclose(cpu_spmv(prob, x), A @ x, rtol=1e-13, atol=1e-13)


@pytest.mark.parametrize("prob", [Problem.poisson2d(16, 12), Problem.poisson3d(8, 7, 6), Problem.stencil27(6, 6, 5)])
def test_cg_solves_manufactured(prob):
    b = prob.rhs("manufactured")
    s = cpu_cg(prob, b, tol=1e-10, maxit=500)
    assert s.converged
    np.testing.assert_allclose(s.x, prob.manufactured_solution(), atol=1e-8)


def test_cg_matches_scipy_direct():
    import scipy.sparse.linalg as spla
    prob = Problem.poisson2d(20, 17)
    b = prob.rhs("random", seed=3)
    s = cpu_cg(prob, b, tol=1e-11, maxit=1000)
    xd = spla.spsolve(prob.to_scipy().tocsc(), b)
    np.testing.assert_allclose(s.x, xd, atol=1e-9)


def test_csr_oracle_equals_stencil_oracle():
    prob = Problem.poisson3d(6, 5, 7)
    b = prob.rhs("random", seed=1)
    s1 = cpu_cg(prob, b, tol=1e-9)
    rowptr, colind, val = prob.to_csr()
    pc = Problem.csr(rowptr, colind, val)
    s2 = cpu_cg(pc, b, tol=1e-9)
    assert s1.iters == s2.iters
    np.testing.assert_allclose(s1.history, s2.history, rtol=1e-9)


def test_abs_vs_rel_tolerance():
    prob = Problem.poisson2d(24)
    b = 100.0 * prob.rhs("ones")
    sa = cpu_cg(prob, b, tol=1e-6, tol_mode="abs")
    sr = cpu_cg(prob, b, tol=1e-6, tol_mode="rel")
    assert sa.rnorm < 1e-6 and sr.rnorm < 1e-6 * sr.r0norm
    assert sr.iters < sa.iters


def test_maxit_falls_through_like_the_reference():
    prob = Problem.poisson2d(32)
    s = cpu_cg(prob, prob.rhs("ones"), tol=1e-14, maxit=5)
    assert s.iters == 5 and not s.converged and len(s.history) == 5


def test_pipelined_variant_converges_to_same_solution():
    prob = Problem.stencil27(8, 8, 8)
    b = prob.rhs("random", seed=5)
    c = cpu_cg(prob, b, tol=1e-9)
    p = cpu_cg(prob, b, tol=1e-9, variant="pipelined")
    assert p.converged and abs(p.iters - c.iters) <= max(2, c.iters // 10)
    np.testing.assert_allclose(p.x, c.x, atol=1e-7)


def test_problem_from_scipy_and_matrix_market(tmp_path):
    import scipy.io
    import scipy.sparse as sp
    A = sp.random(40, 40, density=0.1, random_state=3)
    A = (A + A.T + 40 * sp.identity(40)).tocoo()          # SPD by diagonal dominance
    p1 = Problem.from_scipy(A)
    scipy.io.mmwrite(str(tmp_path / "a.mtx"), A)
    p2 = Problem.from_matrix_market(tmp_path / "a.mtx")
    assert (p1.to_scipy() != p2.to_scipy()).nnz == 0
    b = np.arange(1.0, 41.0)
    s = cpu_cg(p1, b, tol=1e-10)
    assert s.converged
    np.testing.assert_allclose(p1.to_scipy() @ s.x, b, atol=1e-8)


def test_rhs_fill_box_matches_the_per_entry_generator_and_partitions():
    # cgb_rhs_fill_box is what bench.py stages into pinned memory for the end-to-end arm: it must equal the
    # per-entry generator (and hence the device generator) for every rank's box, whatever the decomposition.
    import ctypes as C
    from cuda_mpi_parallel_b200 import Problem, _native as N
    from cuda_mpi_parallel_b200.parallel.partition import gather_global, make_partition
    for prob in (Problem.poisson3d(7, 6, 9), Problem.poisson2d(11, 8), Problem.stencil27(5, 8, 6)):
        g = prob.rhs("random", seed=3)
        ref = np.array([N.lib().cgb_rhs_value(C.byref(prob.stencil), N.RHS_RANDOM, 3, i) for i in range(prob.n)])
        np.testing.assert_array_equal(g, ref)
        for nranks, decomp in ((3, 1), (4, 2 if prob.name != "poisson2d5" else 1)):
            parts = [make_partition(prob, nranks, r, decomp) for r in range(nranks)]
            locs = [prob.rhs_local(p, "random", seed=3) for p in parts]
            np.testing.assert_array_equal(gather_global(prob, parts, locs), g)


def test_csr_constructors_reject_malformed_input_before_touching_the_gpu():
    # A bad offset or column index would make the SpMV read out of bounds on the device (over peer mappings when
    # distributed), so both CSR constructors validate on the host first: this runs without a GPU.
    import ctypes as C
    from cuda_mpi_parallel_b200 import _native as N, make_partition
    L = N.lib()
    good_rp = np.array([0, 2, 3, 5], dtype=np.int32)
    good_ci = np.array([0, 2, 1, 0, 2], dtype=np.int32)
    val = np.ones(5)

    def create(rp, ci, nnz=5, hist_cap=16):
        ctx = C.c_void_p()
        return L.cgb_ctx_create_csr(3, nnz, rp.ctypes.data, ci.ctypes.data, val.ctypes.data, 0, hist_cap, C.byref(ctx))

    cases = {"rowptr[0]": (np.array([1, 2, 3, 5], dtype=np.int32), good_ci, {}),
             "not monotone": (np.array([0, 3, 2, 5], dtype=np.int32), good_ci, {}),
             "does not match nnz": (np.array([0, 2, 3, 4], dtype=np.int32), good_ci, {}),
             "column index": (good_rp, np.array([0, 3, 1, 0, 2], dtype=np.int32), {}),
             "column index ": (good_rp, np.array([0, -1, 1, 0, 2], dtype=np.int32), {}),
             "hist_cap": (good_rp, good_ci, {"hist_cap": 0})}
    for what, (rp, ci, kw) in cases.items():
        rc = create(rp, ci, **kw)
        assert rc < 0, what
        assert what.strip() in L.cgb_last_error().decode(), (what, L.cgb_last_error())
    # the distributed constructor checks columns against the GLOBAL row count and the block against the partition
    prob = Problem.csr(good_rp, good_ci, val)
    part = make_partition(prob, 3, 1)                     # rank 1 owns row 1: one entry
    ctx = C.c_void_p()
    rp1, ci_bad = np.array([0, 1], dtype=np.int32), np.array([3], dtype=np.int32)
    rc = L.cgb_ctx_create_csr_dist(3, C.byref(part), 1, rp1.ctypes.data, ci_bad.ctypes.data, val.ctypes.data, 0, 16, C.byref(ctx))
    assert rc < 0 and "column index" in L.cgb_last_error().decode()
