"""Tier T0: the command-line front end against the reference's literal stdout."""
import json
import re
import subprocess

import numpy as np
import pytest

from cuda_mpi_parallel_b200 import Problem, cpu_cg
from cuda_mpi_parallel_b200.utils import build

REFERENCE_STDOUT = "0.500000\n0.750000\n1.000000\nSuccess\n"   # CUDACG.cu:361-365


@pytest.fixture(scope="module")
def cli():
    build.build_core()
    return str(build.build_cli())


def run(cli, *args):
    return subprocess.run([cli, *args], capture_output=True, text=True, timeout=300)


def test_cpu_variant_reproduces_reference_stdout(cli):
    r = run(cli, "--variant", "cpu")
    assert r.returncode == 0 and r.stdout == REFERENCE_STDOUT and r.stderr == ""


def test_failure_contract_one_line_exit_1(cli):
    r = run(cli, "--problem", "nonsense")
    assert r.returncode == 1 and r.stdout.count("\n") == 1 and r.stderr == ""


def test_baseline_config1_cpu_two_ranks_history_file(cli, tmp_path):
    # BASELINE config 1: 2D 5-point Poisson 64x64, 2 ranks, CPU path.
    h, j, xo = tmp_path / "h.txt", tmp_path / "j.json", tmp_path / "x.txt"
    r = run(cli, "--variant", "cpu", "--problem", "poisson2d5", "--n", "64,64", "--gpus", "2", "--hist-out", str(h),
            "--json-out", str(j), "--x-out", str(xo))
    assert r.returncode == 0 and r.stdout == "Success\n"
    prob = Problem.poisson2d(64, 64)
    ref = cpu_cg(prob, prob.rhs("ones"))
    rows = [l.split() for l in h.read_text().splitlines()]
    assert [int(a) for a, _ in rows] == list(range(1, ref.iters + 1))
    np.testing.assert_allclose([float(b) for _, b in rows], ref.history, rtol=1e-9)
    meta = json.loads(j.read_text())
    assert meta["iters"] == ref.iters and meta["status"] == 1
    np.testing.assert_allclose(np.loadtxt(xo), ref.x, atol=1e-9)


def test_matrix_market_input(cli, tmp_path):
    m = tmp_path / "a.mtx"
    m.write_text("%%MatrixMarket matrix coordinate real symmetric\n% the reference matrix, lower triangle\n"
                 "3 3 4\n1 1 3.0\n3 1 2.0\n2 2 2.0\n3 3 1.0\n")
    r = run(cli, "--variant", "cpu", "--problem", f"csr:{m}")
    assert r.returncode == 0
    got = [float(v) for v in r.stdout.split()[:3]]
    A = np.array([[3, 0, 2], [0, 2, 0], [2, 0, 1.0]])
    np.testing.assert_allclose(got, np.linalg.solve(A, np.ones(3)), atol=1e-5)


@pytest.mark.gpu
def test_gpu_default_run_reproduces_reference_stdout(cli):
    r = run(cli)
    assert r.returncode == 0 and r.stdout == REFERENCE_STDOUT, r.stdout + r.stderr


@pytest.mark.gpu
def test_gpu_stencil_run_matches_oracle(cli, tmp_path):
    h = tmp_path / "h.txt"
    r = run(cli, "--problem", "poisson3d7", "--n", "40,36,44", "--rhs", "random:3", "--hist-out", str(h), "--tol", "1e-8")
    assert r.returncode == 0 and r.stdout == "Success\n", r.stdout
    prob = Problem.poisson3d(40, 36, 44)
    ref = cpu_cg(prob, prob.rhs("random", seed=3), tol=1e-8)
    hist = np.array([float(l.split()[1]) for l in h.read_text().splitlines()])
    assert abs(len(hist) - ref.iters) <= 1
    n = min(len(hist), ref.iters)
    big = ref.history[:n] > 1e-6 * ref.r0norm
    np.testing.assert_allclose(hist[:n][big], ref.history[:n][big], rtol=1e-8)


@pytest.mark.gpu
@pytest.mark.multigpu
def test_gpu_single_process_multi_gpu(cli, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    P = 4 if torch.cuda.device_count() >= 4 else 2
    for variant, decomp in (("fused", "1d"), ("pipelined", "2d" if P == 4 else "1d")):
        h = tmp_path / f"h_{variant}.txt"
        r = run(cli, "--problem", "stencil3d27", "--n", "36,40,44", "--gpus", str(P), "--decomp", decomp, "--variant",
                variant, "--hist-out", str(h), "--tol", "1e-8", "--rhs", "random:4")
        assert r.returncode == 0 and r.stdout == "Success\n", r.stdout + r.stderr
        prob = Problem.stencil27(36, 40, 44)
        ref = cpu_cg(prob, prob.rhs("random", seed=4), tol=1e-8)
        hist = np.array([float(l.split()[1]) for l in h.read_text().splitlines()])
        assert abs(len(hist) - ref.iters) <= 2


@pytest.mark.gpu
@pytest.mark.multigpu
def test_gpu_multi_gpu_restart_from_x_file(cli, tmp_path):
    # checkpoint = x written by --x-out; --x0 restarts from it (r0 = b - A x0) on 2 GPUs incl. the halo of x0
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    common = ["--problem", "poisson3d7", "--n", "40,36,44", "--gpus", "2", "--rhs", "random:6", "--tol", "1e-9"]
    xa, xb, xc = tmp_path / "a.txt", tmp_path / "b.txt", tmp_path / "c.txt"
    ja, jc = tmp_path / "a.json", tmp_path / "c.json"
    assert run(cli, *common, "--x-out", str(xa), "--json-out", str(ja)).returncode == 0
    assert run(cli, *common, "--maxit", "40", "--x-out", str(xb)).returncode == 0
    r = run(cli, *common, "--x0", str(xb), "--x-out", str(xc), "--json-out", str(jc))
    assert r.returncode == 0, r.stdout
    a, c = np.loadtxt(xa), np.loadtxt(xc)
    np.testing.assert_allclose(c, a, atol=1e-7 * np.abs(a).max())
    full, resumed = json.loads(ja.read_text()), json.loads(jc.read_text())
    assert resumed["status"] == 1 and resumed["r0norm"] < full["r0norm"]


def test_python_module_cli_cpu_paths(tmp_path):
    # python -m cuda_mpi_parallel_b200: reference stdout on the CPU path, and BASELINE config 1 over gloo (2 ranks)
    import os
    import sys
    env = dict(os.environ, OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "cuda_mpi_parallel_b200", "--device", "cpu"], capture_output=True, text=True,
                       timeout=300, env=env)
    assert r.returncode == 0 and r.stdout == REFERENCE_STDOUT, r.stdout + r.stderr
    h = tmp_path / "h.txt"
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29911", "-m", "cuda_mpi_parallel_b200", "--problem", "poisson2d5",
                        "--grid", "64", "64", "--device", "cpu", "--hist-out", str(h)], capture_output=True, text=True,
                       timeout=300, env=env)
    assert r.returncode == 0 and "Success" in r.stdout, r.stdout + r.stderr
    prob = Problem.poisson2d(64, 64)
    ref = cpu_cg(prob, prob.rhs("ones"))
    hist = np.array([float(l.split()[1]) for l in h.read_text().splitlines()])
    assert len(hist) == ref.iters
    np.testing.assert_allclose(hist, ref.history, rtol=1e-9)


@pytest.mark.gpu
def test_python_module_cli_gpu(tmp_path):
    import os
    import sys
    j = tmp_path / "j.json"
    r = subprocess.run([sys.executable, "-m", "cuda_mpi_parallel_b200"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout == REFERENCE_STDOUT, r.stdout + r.stderr
    r = subprocess.run([sys.executable, "-m", "cuda_mpi_parallel_b200", "--problem", "poisson3d7", "--n", "48", "--graph", "while",
                        "--json-out", str(j)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout == "Success\n", r.stdout + r.stderr
    meta = json.loads(j.read_text())
    prob = Problem.poisson3d(48)
    assert meta["status"] == 1 and abs(meta["iters"] - cpu_cg(prob, prob.rhs("ones")).iters) <= 1


@pytest.mark.gpu
@pytest.mark.multigpu
def test_gpu_distributed_csr_from_the_cli(cli, tmp_path):
    # `--problem csr:<file> --gpus N`: the general-CSR capability of the reference, row-block distributed over N GPUs
    import scipy.io
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    P = 4 if torch.cuda.device_count() >= 4 else 2
    prob = Problem.poisson3d(14, 12, 10)
    A = prob.to_scipy()
    m = tmp_path / "poisson.mtx"
    scipy.io.mmwrite(str(m), A, symmetry="general")
    xo = tmp_path / "x.txt"
    r = run(cli, "--problem", f"csr:{m}", "--gpus", str(P), "--tol", "1e-9", "--no-print-x", "--x-out", str(xo))
    assert r.returncode == 0 and r.stdout == "Success\n", r.stdout + r.stderr
    x = np.loadtxt(xo)
    np.testing.assert_allclose(A @ x, np.ones(prob.n), atol=1e-7)


def test_python_module_cli_general_csr_cpu(tmp_path):
    # python -m ... --problem csr:<file> / csr-<stencil>: the reference's operand type; several CPU ranks are refused
    import os
    import sys
    import scipy.io
    env = dict(os.environ, OMP_NUM_THREADS="1")
    prob = Problem.poisson3d(9, 8, 7)
    A = prob.to_scipy()
    m, xo = tmp_path / "p.mtx", tmp_path / "x.txt"
    scipy.io.mmwrite(str(m), A, symmetry="general")
    for spec in (f"csr:{m}", "csr-poisson3d7"):
        r = subprocess.run([sys.executable, "-m", "cuda_mpi_parallel_b200", "--device", "cpu", "--problem", spec, "--grid", "9", "8",
                            "7", "--tol", "1e-10", "--x-out", str(xo)], capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0 and r.stdout == "Success\n", r.stdout + r.stderr
        np.testing.assert_allclose(A @ np.loadtxt(xo), np.ones(prob.n), atol=1e-8)
    r = subprocess.run([sys.executable, "-m", "cuda_mpi_parallel_b200", "--device", "cpu", "--problem", "nonsense"],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode != 0 and "unknown problem" in r.stderr
    # two CPU ranks over gloo: row blocks, p all-gathered every iteration
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29915", "-m", "cuda_mpi_parallel_b200", "--device", "cpu", "--problem",
                        f"csr:{m}", "--tol", "1e-10", "--x-out", str(xo)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "Success" in r.stdout, r.stdout + r.stderr
    np.testing.assert_allclose(A @ np.loadtxt(xo), np.ones(prob.n), atol=1e-8)


@pytest.mark.gpu
@pytest.mark.multigpu
def test_python_module_cli_distributed_csr_gpu(tmp_path):
    # the same through torchrun, one rank per GPU: row blocks, remote entries of p read over NVLink
    import os
    import sys
    import scipy.io
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    prob = Problem.poisson3d(14, 12, 10)
    A = prob.to_scipy()
    m, xo = tmp_path / "p.mtx", tmp_path / "x.txt"
    scipy.io.mmwrite(str(m), A, symmetry="general")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29913", "-m", "cuda_mpi_parallel_b200", "--problem", f"csr:{m}",
                        "--tol", "1e-9", "--x-out", str(xo)], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, CGB_TIMEOUT_MS="5000"))
    assert r.returncode == 0 and "Success" in r.stdout, r.stdout + r.stderr
    np.testing.assert_allclose(A @ np.loadtxt(xo), np.ones(prob.n), atol=1e-7)


def test_c_example_compiles_as_c99_and_runs_on_the_host_oracle(tmp_path):
    # include/cgb200.h must stay a valid C header; examples/poisson3d.c is the reference's program flow on the C ABI
    import shutil
    from cuda_mpi_parallel_b200.utils import build
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    build.build_core()
    exe = tmp_path / "poisson3d"
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{build.ROOT / 'include'}",
                        str(build.ROOT / "examples" / "poisson3d.c"), "-o", str(exe), f"-L{build.LIBDIR}", "-lcgb200",
                        f"-Wl,-rpath,{build.LIBDIR}", "-lm"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe), "20", "--cpu"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.endswith("Success\n"), r.stdout + r.stderr
    m = re.search(r"(\d+) iterations, \|\|r\|\| = (\S+) \(recomputed (\S+)\)", r.stdout)
    prob = Problem.poisson3d(20)
    assert int(m.group(1)) == cpu_cg(prob, prob.rhs("ones")).iters
    assert float(m.group(2)) < 1e-7 and float(m.group(3)) < 2e-7
