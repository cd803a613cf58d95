"""`pip install .` / `pip install -e .`: compiles the native libraries for sm_100a with nvcc (the same
`cuda_mpi_parallel_b200.utils.build` that `__graft_entry__.build()` and the first import use) and ships them as
package data. The reference's build recipe is one nvcc line in a comment (CUDACG.cu:1-4)."""
import sys
from pathlib import Path

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = Path(__file__).resolve().parent


class BuildWithNative(build_py):
    def run(self):
        sys.path.insert(0, str(ROOT))
        from cuda_mpi_parallel_b200.utils import build
        build.build_core()
        build.build_baseline()
        super().run()


setup(cmdclass={"build_py": BuildWithNative})
