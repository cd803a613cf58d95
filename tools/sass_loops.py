#!/usr/bin/env python
"""Static size of the inner loops of the band kernels (no GPU needed).

    python tools/sass_loops.py [lib] [kernel-name-regex]

For every kernel whose demangled name matches, lists each backward branch (= loop) that contains an mbarrier
try-wait (the per-plane loops of the TMA band kernels) with its SASS instruction count and opcode mix. Used to
track the per-step instruction overhead of the consumer loops, which is what bounds the 2D and 27-point kernels
(profiles/ROOFLINE.md R2.6).
"""
import collections
import re
import subprocess
import sys


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else "cuda_mpi_parallel_b200/lib/libcgb200.so"
    pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else r"k_p?spmv_tma")
    txt = subprocess.run(f"cuobjdump -sass {lib} | c++filt", shell=True, capture_output=True, text=True).stdout
    name, body = None, []
    funcs = []
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
    for name, body in funcs:
        if not pat.search(name):
            continue
        m = re.search(r"(k_\w+(?:<[^>]*>)?)", name)
        short = m.group(1) if m else name[:60]
        loops = []
        for k, (addr, ins) in enumerate(body):
            m = re.search(r"\bBRA(?:\.[A-Z.]+)?\s+(?:[A-Z0-9!]+,\s*)?`?\(?0x([0-9a-f]+)", ins)
            if m and int(m.group(1), 16) <= addr:
                loops.append((int(m.group(1), 16), addr))
        print(f"{short}: {len(body)} instructions")
        for lo, hi in loops:
            seg = [i for a, i in body if lo <= a <= hi]
            if not any("TRYWAIT" in i for i in seg) or len(seg) < 30:
                continue
            ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", i).split()[0].split(".")[0] for i in seg)
            print(f"   loop {lo:#x}-{hi:#x}: {len(seg)} instructions  " +
                  " ".join(f"{o}:{n}" for o, n in ops.most_common(12)))


if __name__ == "__main__":
    main()
