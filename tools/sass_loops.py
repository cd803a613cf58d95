#!/usr/bin/env python
"""Static size of the inner loops of the band kernels (no GPU needed).

    python tools/sass_loops.py [lib] [kernel-name-regex]

For every kernel whose demangled name matches, lists each backward branch (= loop) that contains an mbarrier
try-wait (the per-plane loops of the TMA band kernels) with its SASS instruction count and opcode mix. Used to
track the per-step instruction overhead of the consumer loops, which is what bounds the 2D and 27-point kernels
(profiles/ROOFLINE.md R2.6). `tests/test_sass_contract.py` pins the sizes as a regression guard.
"""
import collections
import re
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
DEFAULT_LIB = ROOT / "cuda_mpi_parallel_b200" / "lib" / "libcgb200.so"


def cuobjdump_exe():
    return shutil.which("cuobjdump") or ("/usr/local/cuda/bin/cuobjdump" if Path("/usr/local/cuda/bin/cuobjdump").exists() else None)


def functions(lib=DEFAULT_LIB):
    """[(demangled name, [(address, instruction text), ...]), ...] for every kernel in the library."""
    exe = cuobjdump_exe()
    if exe is None:
        raise RuntimeError("cuobjdump not found")
    txt = subprocess.run(f"{exe} -sass {lib} | c++filt", shell=True, capture_output=True, text=True).stdout
    name, body, funcs = None, [], []
    for line in txt.splitlines():
        m = re.match(r"\s*Function : (.*)", line)
        if m:
            if name:
                funcs.append((name, body))
            name, body = m.group(1), []
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if m and name:
            body.append((int(m.group(1), 16), m.group(2).strip()))
    if name:
        funcs.append((name, body))
    return funcs


def kernel_id(name):
    m = re.search(r"(k_\w+(?:<[^>]*>)?)", name)
    return m.group(1) if m else name[:60]


def opcode(ins):
    return re.sub(r"^@!?U?P\d+\s+", "", ins).split()[0].split(".")[0]


def plane_loops(body, lo_size=30, hi_size=450):
    """Innermost-to-outer loops (backward branches) of one kernel that contain an mbarrier try-wait:
    [(first address, last address, [instruction text, ...]), ...]."""
    out = []
    for addr, ins in body:
        m = re.search(r"\bBRA(?:\.[A-Z.]+)?\s+(?:[A-Z0-9!]+,\s*)?`?\(?0x([0-9a-f]+)", ins)
        if m and int(m.group(1), 16) <= addr:
            lo = int(m.group(1), 16)
            seg = [i for a, i in body if lo <= a <= addr]
            if any("TRYWAIT" in i for i in seg) and lo_size <= len(seg) <= hi_size:
                out.append((lo, addr, seg))
    return out


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else DEFAULT_LIB
    pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else r"k_p?spmv_tma")
    for name, body in functions(lib):
        if not pat.search(name):
            continue
        print(f"{kernel_id(name)}: {len(body)} instructions")
        for lo, hi, seg in plane_loops(body, hi_size=10 ** 6):
            ops = collections.Counter(opcode(i) for i in seg)
            print(f"   loop {lo:#x}-{hi:#x}: {len(seg)} instructions  " + " ".join(f"{o}:{n}" for o, n in ops.most_common(12)))


if __name__ == "__main__":
    main()
