"""Tiers T0/T3/T4 on one GPU: the solver against the CPU oracle (x, iteration count, history)."""
import numpy as np
import pytest

from cuda_mpi_parallel_b200 import CGSolver, Problem, cpu_cg

pytestmark = pytest.mark.gpu


def test_ref3x3_matches_reference_answer():
    # CUDACG.cu:79-82 documents x = (0.5, 0.75, 1.0); 3 iterations.
    p = Problem.ref3x3()
    s = CGSolver(p)
    s.set_rhs(p.rhs("default"))
    res = s.solve(tol=1e-7, maxit=2000)
    assert res.status == 1 and res.iters == 3
    np.testing.assert_allclose(s.x().cpu().numpy(), [0.5, 0.75, 1.0], atol=1e-10)
    h = s.history(3)
    np.testing.assert_allclose(h[:2], [8.073855e-01, 1.673428e-01], rtol=1e-6)
    assert h[2] < 1e-7


@pytest.mark.parametrize("prob", [Problem.poisson2d(64, 64), Problem.poisson2d(200, 90), Problem.poisson3d(32, 24, 40), Problem.poisson3d(96, 96, 96),
                                  Problem.stencil27(24, 24, 24)],
                         ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
@pytest.mark.parametrize("use_graph,spmv", [(True, "auto"), (False, "auto"), (True, "simple")])
def test_solver_history_matches_oracle(prob, use_graph, spmv):
    b = prob.rhs("random", seed=7) if prob.n < 500_000 else np.ones(prob.n)
    ref = cpu_cg(prob, b, tol=1e-7, maxit=2000)
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-7, maxit=2000, use_graph=use_graph, chunk=7, spmv=spmv)
    assert res.status == 1
    assert abs(res.iters - ref.iters) <= 1
    n = min(res.iters, ref.iters)
    h = s.history(n)
    big = ref.history[:n] > 1e-6 * ref.r0norm
    np.testing.assert_allclose(h[big], ref.history[:n][big], rtol=1e-8)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-6 * max(1.0, np.abs(ref.x).max()))
    assert res.r0norm == pytest.approx(ref.r0norm, rel=1e-13)


def test_generated_rhs_equals_host_rhs_and_repeat_solves_are_bitwise_identical():
    prob = Problem.poisson3d(48, 40, 36)
    s = CGSolver(prob)
    s.set_rhs("random", seed=5)
    r1 = s.solve(tol=1e-8)
    h1 = s.history(r1.iters)
    x1 = s.x().clone()
    s.set_rhs(prob.rhs("random", seed=5))
    r2 = s.solve(tol=1e-8)
    assert r1.iters == r2.iters
    assert (h1 == s.history(r2.iters)).all()          # bitwise run-to-run
    assert (x1 == s.x()).all()


def test_csr_path_equals_matrix_free_path():
    prob = Problem.poisson3d(20, 18, 16)
    rowptr, colind, val = prob.to_csr()
    b = prob.rhs("random", seed=9)
    s1 = CGSolver(prob)
    s1.set_rhs(b)
    r1 = s1.solve(tol=1e-9)
    s2 = CGSolver(Problem.csr(rowptr, colind, val))
    s2.set_rhs(b)
    r2 = s2.solve(tol=1e-9)
    assert r1.status == r2.status == 1 and abs(r1.iters - r2.iters) <= 1
    np.testing.assert_allclose(s1.x().cpu().numpy(), s2.x().cpu().numpy(), atol=1e-8)


def test_maxit_and_fixed_iters_and_rel_tolerance():
    prob = Problem.poisson2d(256, 256)
    s = CGSolver(prob)
    s.set_rhs("ones")
    r = s.solve(tol=1e-14, maxit=10)
    assert r.status == 2 and r.iters == 10
    r = s.solve(fixed_iters=33, chunk=10, variant="fused")
    assert r.iters == 33 and r.launches == 1 + 3 * 33
    r = s.solve(fixed_iters=33, chunk=10, variant="merged", persistent=False)
    assert r.iters == 33 and r.launches == 1 + 2 * 33
    r = s.solve(fixed_iters=33, chunk=10, variant="merged", persistent=True)   # one cooperative launch per chunk
    assert r.iters == 33 and r.launches == 1 + 4
    r = s.solve(tol=1e-6, tol_mode="rel", maxit=5000)
    assert r.status == 1 and r.rnorm < 1e-6 * r.r0norm


def test_zero_rhs_and_breakdown_are_reported():
    prob = Problem.poisson2d(32, 32)
    s = CGSolver(prob)
    s.set_rhs(np.zeros(prob.n))
    r = s.solve()
    assert r.status == 1 and r.iters == 0
    assert s.x().abs().max().item() == 0.0


@pytest.mark.parametrize("prob", [Problem.stencil27(24, 24, 24), Problem.poisson3d(40, 36, 44), Problem.poisson2d(120, 90)],
                         ids=lambda p: p.name)
def test_pipelined_cg_matches_classic(prob):
    # Ghysels-Vanroose pipelined CG is algebraically classic CG: same x, iteration count within 10 %,
    # history equal to the CPU pipelined oracle up to rounding drift.
    b = prob.rhs("random", seed=13)
    ref_c = cpu_cg(prob, b, tol=1e-8)
    ref_p = cpu_cg(prob, b, tol=1e-8, variant="pipelined")
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-8, variant="pipelined", chunk=6)
    assert res.status == 1
    assert abs(res.iters - ref_c.iters) <= max(2, ref_c.iters // 10)
    assert abs(res.iters - ref_p.iters) <= 2
    n = min(res.iters, ref_p.iters)
    h = s.history(n)
    big = ref_p.history[:n] > 1e-5 * ref_p.r0norm
    np.testing.assert_allclose(h[big], ref_p.history[:n][big], rtol=1e-5)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref_c.x, atol=1e-6)
    r2 = s.solve(tol=1e-14, maxit=7, variant="pipelined", chunk=4)
    assert r2.status == 2 and r2.iters == 7
    assert s.history(7)[6] > 0


@pytest.mark.parametrize("prob", [Problem.poisson3d(40, 36, 44), Problem.poisson3d(130, 37, 21), Problem.stencil27(24, 28, 24),
                                  Problem.poisson3d(512, 24, 20)], ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
def test_recompute_variant_matches_oracle(prob):
    # 72 B/DOF variant: Ap never stored, x/r update re-evaluates the stencil. Same algorithm, same history.
    b = prob.rhs("random", seed=17)
    ref = cpu_cg(prob, b, tol=1e-8)
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-8, variant="recompute", chunk=5)
    assert res.status == 1 and abs(res.iters - ref.iters) <= 1
    n = min(res.iters, ref.iters)
    h = s.history(n)
    big = ref.history[:n] > 1e-6 * ref.r0norm
    np.testing.assert_allclose(h[big], ref.history[:n][big], rtol=1e-8)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-6)
    f = CGSolver(prob)
    f.set_rhs(b)
    rf = f.solve(tol=1e-8, variant="fused", spmv="tma")
    assert rf.iters == res.iters


def test_recompute_variant_rejects_unsupported_problems():
    from cuda_mpi_parallel_b200._native import NativeError
    p = Problem.poisson2d(12, 10)
    rowptr, colind, val = p.to_csr()
    s = CGSolver(Problem.csr(rowptr, colind, val))      # general CSR has no TMA path
    s.set_rhs(np.ones(p.n))
    with pytest.raises(NativeError):
        s.solve(variant="recompute")


@pytest.mark.parametrize("prob", [Problem.poisson3d(40, 36, 44), Problem.poisson3d(130, 37, 21), Problem.stencil27(24, 28, 24),
                                  Problem.poisson3d(512, 24, 20), Problem.poisson3d(384, 20, 9), Problem.poisson3d(33, 5, 3)],
                         ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
@pytest.mark.parametrize("persistent", [True, False], ids=["persistent", "two-kernel"])
def test_merged_variant_matches_oracle(prob, persistent):
    # 64 B/DOF variant: p-update folded into the SpMV (halo rows recomputed in shared memory); either as ONE
    # persistent cooperative kernel per chunk (grid barrier + reduction on the device) or as two kernels per iteration.
    # The persistent kernel sums the per-CTA partials in a different (still fixed) order: equal to rounding, not bitwise.
    b = prob.rhs("random", seed=19)
    ref = cpu_cg(prob, b, tol=1e-8)
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-8, variant="merged", chunk=5, persistent=persistent)
    assert res.status == 1 and abs(res.iters - ref.iters) <= 1
    n = min(res.iters, ref.iters)
    h = s.history(n)
    big = ref.history[:n] > 1e-6 * ref.r0norm
    np.testing.assert_allclose(h[big], ref.history[:n][big], rtol=1e-8)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-6)
    r2 = s.solve(tol=1e-8, variant="merged", chunk=4, use_graph=False, persistent=persistent)     # repeat on the same context
    assert r2.iters == res.iters and (s.history(n) == h).all()
    r3 = s.solve(tol=1e-8, variant="merged", chunk=7, persistent=not persistent)     # the other driver: same arithmetic
    assert r3.iters == res.iters
    np.testing.assert_allclose(s.history(n), h, rtol=1e-12)


@pytest.mark.parametrize("prob", [Problem.poisson3d(64, 64, 64), Problem.poisson2d(300, 200), Problem.stencil27(32, 32, 32)],
                         ids=lambda p: p.name)
def test_persistent_kernel_single_launch_whole_solve(prob):
    # use_graph="while" on the persistent path = ONE kernel launch for the whole solve (plus the init kernel)
    b = prob.rhs("random", seed=3)
    ref = cpu_cg(prob, b, tol=1e-8)
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-8, variant="merged", use_graph="while", persistent=True)
    assert res.status == 1 and abs(res.iters - ref.iters) <= 1 and res.launches == 2
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-6)
    odd = s.solve(fixed_iters=7, variant="merged", chunk=3, persistent=True)    # odd chunks: entry parity alternates
    assert odd.iters == 7
    np.testing.assert_allclose(s.history(7), ref.history[:7], rtol=1e-8)


@pytest.mark.parametrize("variant,prob", [("fused", Problem.poisson2d(150, 110)), ("fused", Problem.stencil27(24, 20, 28)),
                                          ("recompute", Problem.poisson3d(40, 36, 44)), ("merged", Problem.poisson3d(40, 36, 44)),
                                          ("pipelined", Problem.poisson3d(40, 36, 44)), ("fused", Problem.ref3x3())],
                         ids=lambda v: v if isinstance(v, str) else v.name)
def test_while_graph_equals_chunked_graphs(variant, prob):
    # One conditional-WHILE graph per solve: same iterations and bitwise the same history as the chunked driver.
    b = prob.rhs("default") if prob.name == "ref3x3" else prob.rhs("random", seed=23)
    s = CGSolver(prob)
    s.set_rhs(b)
    r1 = s.solve(tol=1e-8, variant=variant, use_graph=True, chunk=6)
    h1 = s.history(r1.iters)
    x1 = s.x().clone()
    r2 = s.solve(tol=1e-8, variant=variant, use_graph="while")
    assert r2.status == r1.status == 1 and r2.iters == r1.iters
    assert (s.history(r2.iters) == h1).all() and (s.x() == x1).all()
    if prob.n > 3:                                   # (the 3x3 system is solved exactly before 9 iterations)
        r3 = s.solve(tol=1e-30, maxit=9, variant=variant, use_graph="while")
        assert r3.status == 2 and r3.iters == 9
    s.set_rhs(np.zeros(prob.n))                      # done is already set at graph entry: the loop must still end
    r4 = s.solve(variant=variant, use_graph="while")
    assert r4.iters == 0


@pytest.mark.parametrize("variant,prob", [("fused", Problem.poisson2d(96, 80)), ("merged", Problem.poisson3d(40, 36, 44)),
                                          ("pipelined", Problem.stencil27(20, 24, 22)), ("fused", Problem.ref3x3())],
                         ids=lambda v: v if isinstance(v, str) else v.name)
def test_warm_start_and_checkpoint_restart(variant, prob, tmp_path):
    # The reference always starts from x0 = 0 (CUDACG.cu:177-179,247). Warm start: r0 = b - A x0.
    b = prob.rhs("default") if prob.name == "ref3x3" else prob.rhs("random", seed=29)
    A = prob.to_scipy()
    s = CGSolver(prob)
    s.set_rhs(b)
    full = s.solve(tol=1e-9, variant=variant)
    x_full = s.x().cpu().numpy()
    # interrupted run: a few iterations, checkpoint, "new process", resume
    part = s.solve(tol=1e-9, variant=variant, maxit=max(1, full.iters // 2))
    s.save_checkpoint(tmp_path / "ck.npz", iters_done=part.iters)
    x_part = s.x().cpu().numpy()
    t = CGSolver(prob)
    t.set_rhs(b)
    assert t.load_checkpoint(tmp_path / "ck.npz") == part.iters
    res = t.solve(tol=1e-9, variant=variant, warm_start=True)
    assert res.status == 1
    assert res.r0norm == pytest.approx(np.linalg.norm(b - A @ x_part), rel=1e-9)      # r0 = b - A x0
    np.testing.assert_allclose(t.x().cpu().numpy(), x_full, atol=1e-7 * max(1.0, np.abs(x_full).max()))
    # warm start from the converged solution needs no iterations worth mentioning
    res2 = t.solve(tol=1e-6, variant=variant, warm_start=True)
    assert res2.iters <= 1


def test_host_history_streams_every_residual_to_pinned_memory():
    import torch
    prob = Problem.poisson3d(40, 36, 44)
    s = CGSolver(prob)
    s.set_rhs("random", seed=31)
    hist = torch.zeros(600, dtype=torch.float64).pin_memory()
    s.set_host_history(hist)
    for variant in ("merged", "pipelined", "fused"):
        hist.zero_()
        res = s.solve(tol=1e-8, variant=variant)
        assert res.status == 1
        np.testing.assert_array_equal(hist[:res.iters].numpy(), s.history(res.iters))
        assert float(hist[res.iters]) == 0.0
    s.set_host_history(None)
    hist.zero_()
    s.solve(tol=1e-8)
    assert float(hist.abs().max()) == 0.0


@pytest.mark.parametrize("variant", ["fused", "recompute", "merged", "pipelined"])
@pytest.mark.parametrize("prob", [Problem.poisson3d(1030, 9, 7), Problem.stencil27(1537, 6, 5)],
                         ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
def test_x_tiled_rows_all_variants(prob, variant):
    # nx > 512: the TMA kernels cut rows into 512-cell tiles whose x halo arrives in 2-column boxes.
    b = prob.rhs("random", seed=37)
    ref = cpu_cg(prob, b, tol=1e-8, variant="pipelined" if variant == "pipelined" else "classic")
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-8, variant=variant, spmv="tma", chunk=5)
    assert res.status == 1 and abs(res.iters - ref.iters) <= 2
    n = min(res.iters, ref.iters)
    h = s.history(n)
    big = ref.history[:n] > 1e-5 * ref.r0norm
    np.testing.assert_allclose(h[big], ref.history[:n][big], rtol=1e-5 if variant == "pipelined" else 1e-8)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-6)


@pytest.mark.parametrize("variant", ["fused", "recompute", "merged", "pipelined"])
@pytest.mark.parametrize("prob", [Problem.poisson2d(200, 90), Problem.poisson2d(33, 17), Problem.poisson2d(2300, 12),
                                  Problem.poisson2d(512, 40), Problem.poisson2d(2304, 24)],
                         ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
def test_2d_problems_on_the_tma_kernels(prob, variant):
    # a 2D grid is a one-row-per-plane 3D grid: same band kernels, no y taps
    b = prob.rhs("random", seed=41)
    ref = cpu_cg(prob, b, tol=1e-8, variant="pipelined" if variant == "pipelined" else "classic")
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(tol=1e-8, variant=variant, spmv="tma", chunk=5)
    assert res.status == 1 and abs(res.iters - ref.iters) <= 2
    n = min(res.iters, ref.iters)
    h = s.history(n)
    big = ref.history[:n] > 1e-5 * ref.r0norm
    np.testing.assert_allclose(h[big], ref.history[:n][big], rtol=1e-5 if variant == "pipelined" else 1e-8)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-6 * max(1.0, np.abs(ref.x).max()))


@pytest.mark.parametrize("variant", ["fused", "recompute", "merged"])
@pytest.mark.parametrize("prob", [Problem.poisson3d(32, 1280, 24), Problem.stencil27(32, 1280, 24)],
                         ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
def test_several_items_per_cta(prob, variant):
    # 80 bands x 3 z chunks = 240 (band, chunk) items on at most 148 CTAs: CTAs run one or two items back to back
    # through the same stage ring (tests/test_tma_geometry.py pins the decomposition)
    b = prob.rhs("random", seed=13)
    K = 40
    ref = cpu_cg(prob, b, tol=1e-30, maxit=K)
    s = CGSolver(prob)
    s.set_rhs(b)
    res = s.solve(fixed_iters=K, variant=variant, spmv="tma", chunk=9)
    assert res.iters == K and res.status in (0, 2)
    np.testing.assert_allclose(s.history(K), ref.history[:K], rtol=1e-8)
    np.testing.assert_allclose(s.x().cpu().numpy(), ref.x, atol=1e-8 * max(1.0, np.abs(ref.x).max()))


def test_iterate_streamed_copies_each_residual_to_pinned_memory():
    import torch
    prob = Problem.poisson3d(40, 36, 44)
    s = CGSolver(prob)
    s.set_rhs("random", seed=43)
    hist = torch.zeros(12, dtype=torch.float64).pin_memory()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        s.begin(fixed_iters=12, stream=st)
        s.iterate_streamed(12, hist)
        st.synchronize()
        res = s.finish()
    assert res.iters == 12
    np.testing.assert_array_equal(hist.numpy(), s.history(12))


def test_poll_without_wait_never_blocks_and_catches_up():
    # cgb_solve_poll(wait=0) returns the newest snapshot whose copy has landed and never synchronises the stream
    import ctypes as C
    import time
    from cuda_mpi_parallel_b200 import _native as N
    prob = Problem.poisson3d(96, 96, 96)
    s = CGSolver(prob)
    s.set_rhs("random", seed=5)
    s.begin(fixed_iters=300, chunk=50)
    s.iterate(300)
    r = N.SolveResult()
    seen = []
    for _ in range(2000):
        N.check(N.lib().cgb_solve_poll(s._ctx, 0, C.byref(r)))
        seen.append(r.iters)
        if r.iters == 300:
            break
        time.sleep(0.001)
    assert seen == sorted(seen) and seen[-1] == 300, seen[-5:]
    res = s.finish()
    assert res.iters == 300
