"""Multi-process tiers: gloo world_size 2/4 on the CPU (runs anywhere); nccl + cudaIpc fabric on >= 2 GPUs."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def _run(nproc, args, port, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / "mp_worker.py"), *args]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
    lines = [l for l in r.stdout.splitlines() if l.startswith("MPRESULT ")]
    assert r.returncode == 0 and lines, f"worker failed rc={r.returncode}\n{r.stdout[-2000:]}\n{r.stderr[-3000:]}"
    return json.loads(lines[-1][len("MPRESULT "):])


@pytest.mark.parametrize("nproc,name,dims,decomp", [(2, "poisson2d5", "64x64", 1), (4, "poisson3d7", "10x12x16", 2),
                                                    (4, "stencil3d27", "8x10x12", 2), (3, "csr-stencil3d27", "7x6x5", 1),
                                                    (2, "csr-poisson2d5", "31x17", 1)])
def test_cpu_distributed_cg_gloo(nproc, name, dims, decomp):
    # BASELINE config 1: 2D 5-pt 64x64, 2 ranks, CPU path. csr-*: general CSR in (uneven) row blocks, p all-gathered.
    res = _run(nproc, ["cpu", name, dims, str(decomp)], port=29600 + nproc + decomp)
    assert res["ok"] and res["ranks_identical"], res


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("name,dims,decomp", [("poisson3d7", "64x48x40", 1), ("poisson2d5", "96x70", 1),
                                              ("stencil3d27", "40x36x32", 1), ("poisson3d7", "96x50x44", 2),
                                              ("stencil3d27", "36x42x38", 2), ("poisson3d7", "512x64x48", 1)])
def test_gpu_fabric_torchrun(name, dims, decomp):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 2 if n < 4 else 4
    res = _run(nproc, ["gpu", name, dims, str(decomp)], port=29700 + decomp, timeout=240)
    assert res["ok"] and res["ranks_identical"], res


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("name,dims,decomp", [("stencil3d27", "40x36x32", 1), ("poisson3d7", "96x50x44", 2),
                                              ("poisson2d5", "96x70", 1)])
def test_gpu_fabric_pipelined_cg(name, dims, decomp):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 2 if n < 4 else 4
    res = _run(nproc, ["gpu", name, dims, str(decomp), "pipelined"], port=29720 + decomp, timeout=240)
    assert res["ok"] and res["ranks_identical"], res


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("name,dims,decomp", [("poisson3d7", "64x48x40", 1), ("stencil3d27", "36x42x38", 2)])
def test_gpu_fabric_recompute_variant(name, dims, decomp):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 2 if n < 4 else 4
    res = _run(nproc, ["gpu", name, dims, str(decomp), "recompute"], port=29740 + decomp, timeout=240)
    assert res["ok"] and res["ranks_identical"], res


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("name,dims", [("poisson3d7", "64x48x40"), ("stencil3d27", "36x42x38"), ("poisson3d7", "512x20x9"),
                                       ("poisson3d7", "1030x6x12"), ("poisson2d5", "300x40")])
def test_gpu_fabric_merged_variant(name, dims):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 2 if n < 4 else 4
    res = _run(nproc, ["gpu", name, dims, "1", "merged"], port=29760, timeout=240)
    assert res["ok"] and res["ranks_identical"], res


@pytest.mark.gpu
@pytest.mark.multigpu
def test_gpu_fabric_soak_thousands_of_handshakes():
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 2 if n < 4 else 4
    res = _run(nproc, ["soak", "poisson3d7", "32x16x24", "1", "merged"], port=29780, timeout=300)
    assert res["ok"], res


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("name,dims", [("csr-stencil3d27", "13x11x9"), ("csr-poisson2d5", "70x45")])
def test_gpu_distributed_general_csr(name, dims):
    # beyond the reference (single GPU): row-block CSR whose SpMV loads remote entries of p over NVLink
    n = _ngpu()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    nproc = 2 if n < 4 else 4
    res = _run(nproc, ["gpu", name, dims, "1", "fused"], port=29790, timeout=240)
    assert res["ok"] and res["ranks_identical"], res


# ---- tiers that use every GPU of the box (8 on the B200 node): T3 partition invariance at full width ----

def _all_gpus(minimum):
    n = _ngpu()
    if n < minimum:
        pytest.skip(f"needs >= {minimum} GPUs")
    return min(n, 8)


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("name,dims,decomp,variant", [
    ("poisson3d7", "64x48x64", 1, "merged"), ("poisson3d7", "64x48x64", 1, "merged+persist"),
    ("stencil3d27", "48x40x64", 1, "merged"), ("stencil3d27", "48x40x64", 1, "merged+persist"),
    ("poisson2d5", "200x160", 1, "merged+persist"), ("poisson3d7", "1030x8x64", 1, "merged+persist"),
    ("poisson2d5", "1280x160", 1, "merged"), ("poisson2d5", "512x128", 1, "merged+persist"),
    ("poisson3d7", "48x64x64", 2, "recompute"), ("stencil3d27", "40x64x64", 2, "fused"),
    ("stencil3d27", "40x64x64", 2, "pipelined"), ("poisson3d7", "64x48x64", 1, "fused+while"),
    ("poisson3d7", "64x48x64", 1, "merged+while")])
def test_gpu_fabric_all_gpus(name, dims, decomp, variant):
    # every visible GPU (8 ranks on the B200 box): z slabs, 2x4 pencils (decomp=2), the persistent kernel, and the
    # conditional-WHILE graph at N > 1
    nproc = _all_gpus(2)
    res = _run(nproc, ["gpu", name, dims, str(decomp), variant], port=29810 + decomp, timeout=300)
    assert res["ok"] and res["ranks_identical"] and res["world"] == nproc, res


@pytest.mark.gpu
@pytest.mark.multigpu
@pytest.mark.parametrize("variant", ["merged", "merged+persist"])
def test_gpu_fabric_soak_all_gpus(variant):
    nproc = _all_gpus(2)
    res = _run(nproc, ["soak", "poisson3d7", "32x16x64", "1", variant], port=29830, timeout=300)
    assert res["ok"] and res["world"] == nproc, res
