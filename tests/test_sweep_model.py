"""Executable statement of the band kernels' z sweep (csrc/kernels/spmv_tma.cu: Run2, plane_ct, run_advance, emit).

With the in-plane sums of plane k - S0 (the cell), S1 (its 4 in-plane face neighbours; 2 on a 2D grid), S2 (its 4
in-plane diagonal neighbours) - define
    C_k = cc*S0_k + cf*S1_k + ce*S2_k      plane k's contribution to output plane k
    T_k = cf*S0_k + ce*S1_k + cn*S2_k      plane k's contribution to output planes k-1 and k+1
so that (A p)_k = C_k + T_{k-1} + T_{k+1}. A sweep over a z chunk [k0, k1) loads planes k0-1 .. k1 (two more than it
emits), carries part = C_k + T_{k-1} and T_k, and emits plane k when plane k+1 has arrived; a chunk next to a ghost
plane sweeps towards it (downwards for the bottom chunk) with the same recurrences mirrored. This numpy model is
checked against the CPU oracle, which replaces cusparseSpMV (CUDACG.cu:288-291) for the structured operators.
"""
import numpy as np
import pytest

from cuda_mpi_parallel_b200 import Problem, cpu_spmv


def _plane_sums(P, has_y):
    """S0, S1, S2 of one plane P[y, x] with zero (Dirichlet) padding."""
    Z = np.pad(P, 1)
    c = Z[1:-1, 1:-1]
    s1 = Z[1:-1, :-2] + Z[1:-1, 2:]
    s2 = np.zeros_like(c)
    if has_y:
        s1 = s1 + Z[:-2, 1:-1] + Z[2:, 1:-1]
        s2 = Z[:-2, :-2] + Z[:-2, 2:] + Z[2:, :-2] + Z[2:, 2:]
    return c, s1, s2


def sweep_model(prob, p, zs):
    st = prob.stencil
    cc, cf = st.c_center, st.c_face
    ce, cn = (st.c_edge, st.c_corner) if prob.name == "stencil3d27" else (0.0, 0.0)
    gx, gy, gz = prob.dims
    has_y = prob.name != "poisson2d5"                 # a 2D grid is stored one row per plane: no y taps
    V = p.reshape(gz, gy, gx)
    out = np.zeros_like(V)
    zchunk = -(-gz // zs)

    def CT(q):
        if q < 0 or q >= gz:                          # outside the domain: zero plane
            z = np.zeros((gy, gx))
            return z, z
        s0, s1, s2 = _plane_sums(V[q], has_y)
        return cc * s0 + cf * s1 + ce * s2, cf * s0 + ce * s1 + cn * s2

    for c in range(-(-gz // zchunk)):
        k0, k1 = c * zchunk, min(gz, (c + 1) * zchunk)
        down = c == 0 and zs > 1 and c % 2 == 0       # (any chunk may sweep either way; exercise both)
        order = range(k1, k0 - 2, -1) if down else range(k0 - 1, k1 + 1)
        part = t_prev = None
        for s, q in enumerate(order):                 # step s loads plane q
            C, T = CT(q)
            if s >= 2:                                # plane loaded at step s-1 is complete: part + T of its other side
                out[q + 1 if down else q - 1] = part + T
            part = C + (t_prev if t_prev is not None else 0.0)
            t_prev = T
    return out.reshape(-1)


@pytest.mark.parametrize("prob", [Problem.poisson2d(13, 9), Problem.poisson3d(7, 6, 5), Problem.stencil27(6, 5, 7),
                                  Problem.stencil27(9, 8, 16)], ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
@pytest.mark.parametrize("zs", [1, 2, 3])
def test_partial_sum_sweep_equals_the_operator(prob, zs):
    rng = np.random.default_rng(3)
    p = rng.standard_normal(prob.n)
    want = cpu_spmv(prob, p)
    got = sweep_model(prob, p, zs)
    assert np.abs(got - want).max() <= 1e-13 * np.abs(want).max()
