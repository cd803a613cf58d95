"""What the compiled sm_100a code must contain, checked on the CPU with cuobjdump (no GPU needed).

* the band kernels move their tiles with TMA (`UTMALDG` = cp.async.bulk.tensor) and mbarriers (`SYNCS`);
* the L1-path SpMV never loads the exchanged vector through the non-coherent path: peers write its ghost layers
  while the kernel runs, and an acquire does not order `ld.global.nc` (`LDG...CONSTANT`);
* the per-plane loops of the band kernels stay lean and (almost) free of local-memory traffic - they are
  issue-bound on 2D and 27-point grids (profiles/ROOFLINE.md R2.6), so a silent 30 % growth is a performance bug.
"""
import re
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tools"))
import sass_loops  # noqa: E402


@pytest.fixture(scope="module")
def kernels():
    if sass_loops.cuobjdump_exe() is None:
        pytest.skip("cuobjdump not available")
    from cuda_mpi_parallel_b200.utils import build
    build.build_core()
    return {sass_loops.kernel_id(n): body for n, body in sass_loops.functions()}


def _count(body, pat):
    return sum(1 for _, i in body if re.search(pat, i))


def test_band_kernels_use_tma_and_mbarriers(kernels):
    names = [k for k in kernels if re.match(r"k_p?spmv_tma<|k_cg_persist<", k)]
    assert len(names) == 15, names                       # 3 merged + 9 SpMV modes + 3 persistent
    for k in names:
        assert _count(kernels[k], r"\bUTMALDG") >= 10, k
        assert _count(kernels[k], r"\bSYNCS") >= 16, k
    # and nothing GEMM-shaped exists in this solver: no tensor-core instructions to look for
    assert not any(_count(b, r"\bUTC\w*MMA|\bHMMA|\bDMMA") for b in kernels.values())


def test_exchanged_vector_is_never_loaded_non_coherently(kernels):
    for k in ("k_spmv<1>", "k_spmv<2>", "k_spmv<3>"):
        assert k in kernels
        assert _count(kernels[k], r"LDG[.\w]*\.CONSTANT") == 0, k


def test_peer_traffic_is_inside_the_compute_kernels(kernels):
    # halo pushes and the all-reduce are system-scope stores/loads issued by the kernels themselves
    for k in ("k_pspmv_tma<2>", "k_spmv_tma<2, 2>", "k_pupdate", "k_init"):
        assert _count(kernels[k], r"\.SYS\b") >= 4, k


# (kernel, static size of its innermost plane loop when this was written); sizes include ~50 instructions of the
# never-taken time-out path of the bounded mbarrier wait
LOOPS = [("k_pspmv_tma<1>", 169), ("k_spmv_tma<1, 2>", 163), ("k_pspmv_tma<2>", 266), ("k_spmv_tma<2, 2>", 243),
         ("k_pspmv_tma<3>", 328), ("k_spmv_tma<3, 2>", 319), ("k_spmv_tma<2, 0>", 187), ("k_spmv_tma<3, 0>", 261)]


@pytest.mark.parametrize("kernel,size", LOOPS)
def test_plane_loops_stay_lean(kernels, kernel, size):
    loops = sass_loops.plane_loops(kernels[kernel])
    assert loops, f"no plane loop found in {kernel}"
    inner = min(loops, key=lambda l: len(l[2]))[2]
    assert len(inner) <= size * 1.10, f"{kernel}: plane loop grew from {size} to {len(inner)} instructions"
    assert sum(1 for i in inner if re.search(r"\b(STL|LDL)\b", i)) <= 4, f"{kernel}: spills inside the plane loop"
    # no integer division by the run-time stage count in the loop (it used to cost ~25 instructions per plane)
    assert not any(re.search(r"\bI2F|\bMUFU\.RCP", i) for i in inner), kernel
