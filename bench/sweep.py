#!/usr/bin/env python
"""Strong/weak-scaling sweep (BASELINE.json config 5): runs bench.py for one GPU count over several grids.

    python bench/sweep.py --gpus 8 --grids 256 512 1024 --out profiles/sweep_n8.jsonl
    python bench/sweep.py --gpus 8 --weak 512 512 64 --out profiles/sweep_weak_n8.jsonl   # per-GPU box fixed

One GPU count per call because a `gpurun` box has a fixed number of GPUs (and is charged per GPU);
`python bench/report.py profiles/sweep_*.jsonl` merges the files into a scaling table.
"""
import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def run_bench(n, extra, port):
    if n == 1:
        cmd = [sys.executable, str(ROOT / "bench.py"), "--gpus", "1"]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
               "127.0.0.1", "--master-port", str(port), str(ROOT / "bench.py"), "--gpus", str(n)]
    r = subprocess.run(cmd + extra, capture_output=True, text=True, timeout=600, cwd=ROOT)
    for line in reversed(r.stdout.splitlines()):
        if line.startswith("{"):
            return json.loads(line)
    return {"error": (r.stdout + r.stderr)[-800:]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, required=True)
    ap.add_argument("--grids", type=int, nargs="*", default=[])
    ap.add_argument("--weak", type=int, nargs=3, default=None, help="per-GPU box nx ny nz (z grows with the GPU count)")
    ap.add_argument("--problem", default="poisson3d7")
    ap.add_argument("--variant", default="auto")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--out", required=True)
    a = ap.parse_args()
    jobs = [("strong", [g, g, g]) for g in a.grids]
    if a.weak:
        jobs.append(("weak", [a.weak[0], a.weak[1], a.weak[2] * a.gpus]))
    out = Path(a.out)
    out.parent.mkdir(parents=True, exist_ok=True)
    with out.open("a") as f:
        for k, (mode, grid) in enumerate(jobs):
            extra = ["--steps", str(a.steps), "--warmup", str(a.warmup), "--no-e2e", "--problem", a.problem, "--variant",
                     a.variant, "--grid", *map(str, grid)]
            res = run_bench(a.gpus, extra, 29800 + k)
            res.update({"sweep_mode": mode, "grid": grid, "n_gpus": a.gpus})
            f.write(json.dumps(res) + "\n")
            f.flush()
            print(mode, grid, a.gpus, res.get("value"), res.get("ms_per_step"), res.get("error", ""), flush=True)


if __name__ == "__main__":
    main()
