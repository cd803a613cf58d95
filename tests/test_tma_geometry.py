"""Host-side checks of the TMA band-kernel work decomposition (no GPU needed)."""
import ctypes as C

import pytest

from cuda_mpi_parallel_b200 import Problem, _native as N, make_layout, make_partition

NAMES = ["LX", "LY", "RB", "HY", "prow", "NBX", "ntx", "nby", "nbands", "zs", "zchunk", "nitems", "stage_bytes",
         "nstages", "grid", "nhalo"]


def geo(prob, nranks=1, rank=0, mode=0, nsm=148):
    part = make_partition(prob, nranks, rank)
    lay = make_layout(prob, part)
    out = (C.c_int * 16)()
    lo, hi = int(part.nbr[1] >= 0), int(part.nbr[7] >= 0)       # directions (0,-1) and (0,+1)
    ok = N.lib().cgb_debug_tma_geometry(C.byref(prob.stencil), C.byref(lay), lo, hi, nsm, mode, out)
    return (dict(zip(NAMES, out)) if ok else None), lay


@pytest.mark.parametrize("prob", [Problem.poisson3d(512), Problem.poisson3d(384), Problem.poisson3d(64, 48, 40),
                                  Problem.poisson3d(1030, 9, 7), Problem.stencil27(1537, 6, 5), Problem.poisson2d(8192),
                                  Problem.poisson2d(33, 17), Problem.poisson2d(2300, 12), Problem.poisson3d(7, 5, 3)],
                         ids=lambda p: f"{p.name}-{'x'.join(map(str, p.dims))}")
@pytest.mark.parametrize("mode", [0, 4, 5])
def test_bands_tiles_and_chunks_cover_the_box(prob, mode):
    g, lay = geo(prob, mode=mode)
    assert g is not None
    tile = 2 * g["LX"]
    assert g["ntx"] * tile >= lay.nx and (g["ntx"] - 1) * tile < lay.nx          # x tiles cover the row
    assert g["nby"] * g["RB"] >= lay.ny and (g["nby"] - 1) * g["RB"] < lay.ny    # bands cover the rows
    assert g["zs"] * g["zchunk"] >= lay.nz and (g["zs"] - 1) * g["zchunk"] < lay.nz
    assert g["nitems"] == g["nbands"] * g["zs"] == g["ntx"] * g["nby"] * g["zs"]
    assert g["prow"] == g["RB"] + 2 * g["HY"]
    assert 2 <= g["nstages"] <= 8 and g["nstages"] * g["stage_bytes"] <= 200 * 1024   # fits one CTA's shared memory
    assert 1 <= g["grid"] <= 148 and g["LX"] * g["LY"] <= 512                          # one CTA per SM, 512 consumers
    assert (g["HY"], g["RB"]) == ((0, 1) if prob.name == "poisson2d5" else (1, 2 * g["LY"]))


def test_headline_geometry_and_halo_ctas():
    prob = Problem.poisson3d(512)
    g, _ = geo(prob, mode=5)
    assert (g["RB"], g["nbands"], g["zs"], g["grid"], g["nhalo"]) == (4, 128, 1, 128, 0)
    # a middle rank of the 8-GPU run: two half sweeps per band, 20 idle SMs push the halo
    g, lay = geo(prob, nranks=8, rank=3, mode=5)
    assert lay.nz == 64 and (g["zs"], g["zchunk"], g["nitems"]) == (2, 32, 256)
    assert g["grid"] == 148 and g["nhalo"] == 20
    # an end rank sweeps once, towards its only ghost plane
    g, _ = geo(prob, nranks=8, rank=0, mode=5)
    assert g["zs"] == 1 and g["nhalo"] == 20
    # the plain SpMV / update kernels never need halo CTAs
    assert geo(prob, nranks=8, rank=3, mode=4)[0]["nhalo"] == 0


def test_z_split_uses_all_sms_when_one_round_would_leave_them_idle():
    # 384 rows = 96 bands: one round would occupy 96 of 148 SMs; three z chunks give 288 items = 2 rounds on 148 CTAs
    for prob in (Problem.stencil27(384, 384, 384), Problem.poisson3d(384)):
        for mode in (0, 4, 5):
            g, _ = geo(prob, mode=mode)
            assert (g["nbands"], g["zs"], g["zchunk"], g["nitems"], g["grid"]) == (96, 3, 128, 288, 148)
    # 1024^3: 512 bands = 3.46 rounds; two z chunks make it 6.92
    g, _ = geo(Problem.poisson3d(1024), mode=5)
    assert (g["nbands"], g["zs"], g["grid"]) == (512, 2, 148)
    # a small tall grid used by the GPU tests to run several items per CTA
    g, _ = geo(Problem.poisson3d(32, 1280, 24), mode=5)
    assert (g["nbands"], g["zs"], g["nitems"], g["grid"]) == (80, 3, 240, 148)
    # the headline grid keeps its single round (measured: a finer split is 7-10 % slower there)
    for mode in (0, 4, 5):
        g, _ = geo(Problem.poisson3d(512), mode=mode)
        assert (g["zs"], g["grid"]) == (1, 128)


def test_csr_has_no_tma_path():
    p = Problem.poisson2d(12, 10)
    rowptr, colind, val = p.to_csr()
    assert geo(Problem.csr(rowptr, colind, val))[0] is None


# ---- property test: any grid / rank count / mode gives a decomposition that covers the box and fits the chip ----
from hypothesis import given, settings, strategies as st  # noqa: E402


@st.composite
def _geo_cases(draw):
    kind = draw(st.sampled_from(["poisson2d5", "poisson3d7", "stencil3d27"]))
    nx = draw(st.sampled_from([1, 7, 33, 256, 384, 512, 1030, 2304, 4096]))
    ny = draw(st.integers(1, 1500))
    nz = draw(st.integers(1, 1100))
    nranks = draw(st.sampled_from([1, 2, 3, 8]))
    return kind, nx, ny, nz, nranks, draw(st.integers(0, 7)), draw(st.sampled_from([0, 1, 4, 5])), draw(st.sampled_from([147, 148]))


@given(_geo_cases())
@settings(max_examples=300, deadline=None)
def test_geometry_properties(case):
    kind, nx, ny, nz, nranks, rank, mode, nsm = case
    prob = Problem.from_name(kind, [nx, ny, nz])
    rank %= nranks
    try:
        g, lay = geo(prob, nranks=nranks, rank=rank, mode=mode, nsm=nsm)
    except Exception:
        return                                            # more ranks than planes: the partition is refused
    if g is None:
        return                                            # (no stage fits: the solver falls back to the L1-path kernels)
    tile = 2 * g["LX"]
    assert g["ntx"] * tile >= lay.nx and (g["ntx"] - 1) * tile < lay.nx
    assert g["nby"] * g["RB"] >= lay.ny and (g["nby"] - 1) * g["RB"] < lay.ny
    assert g["zs"] * g["zchunk"] >= lay.nz and (g["zs"] - 1) * g["zchunk"] < lay.nz and g["zchunk"] >= 1
    assert g["nitems"] == g["ntx"] * g["nby"] * g["zs"]
    assert 1 <= g["grid"] <= nsm and 0 <= g["nhalo"] <= 24
    sweepers = g["grid"] - g["nhalo"]
    assert 1 <= sweepers <= g["nitems"]                   # no sweeping CTA without an item
    assert 2 <= g["nstages"] <= 8 and g["nstages"] * g["stage_bytes"] <= 200 * 1024
    # the chosen split is never predicted slower than one item per CTA (rounds x (planes per item + 2))
    rounds = -(-g["nitems"] // sweepers) if g["nhalo"] == 0 else -(-g["nitems"] // sweepers)
    one = max(1, min(nsm // g["nbands"], max(1, lay.nz // 8)))
    zc1 = -(-lay.nz // one)
    one_cost = -(-(g["nbands"] * -(-lay.nz // zc1)) // nsm) * (zc1 + 2)
    if g["nhalo"] == 0 and not (mode == 5 and g["zs"] == 2 and one == 1):   # (two half sweeps of a middle rank: forced)
        assert rounds * (g["zchunk"] + 2) <= one_cost
