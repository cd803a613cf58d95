"""Build the UNMODIFIED reference program for the B0 check (stdout == the four documented lines).

`/root/reference/CUDACG.cu` is compiled as-is; the only addition is a command-line macro that maps the
cuSPARSE 11 name it uses (CUSPARSE_MV_ALG_DEFAULT, CUDACG.cu:275,291 — removed in cuSPARSE 12) to its
cuSPARSE 12 replacement.  Output goes to baseline/_ref/ (git-ignored, shipped to the GPU box by gpurun).
The source itself is never copied into this repository.
"""
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
SRC = Path("/root/reference/CUDACG.cu")
OUT = ROOT / "baseline" / "_ref" / "ref_cudacg"


def build() -> Path | None:
    if not SRC.exists():
        return OUT if OUT.exists() else None
    OUT.parent.mkdir(parents=True, exist_ok=True)
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-DCUSPARSE_MV_ALG_DEFAULT=CUSPARSE_SPMV_ALG_DEFAULT",
           str(SRC), "-o", str(OUT), "-lcublas", "-lcusparse"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        print(r.stdout + r.stderr, file=sys.stderr)
        return None
    return OUT


if __name__ == "__main__":
    print(build())
