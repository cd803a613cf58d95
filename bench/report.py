#!/usr/bin/env python
"""Merge sweep_*.jsonl files into a markdown scaling table (efficiency relative to the smallest N of a grid)."""
import json
import sys
from collections import defaultdict
from pathlib import Path


def main():
    rows = []
    for p in sys.argv[1:]:
        for line in Path(p).read_text().splitlines():
            if line.strip():
                rows.append(json.loads(line))
    by = defaultdict(dict)
    for r in rows:
        if "value" in r:
            key = (r["sweep_mode"], tuple(r["grid"]) if r["sweep_mode"] == "strong" else tuple(r["grid"][:2]) + (r["grid"][2] // r["n_gpus"],))
            by[key][r["n_gpus"]] = r
    print("| mode | grid (strong: global, weak: per GPU) | GPUs | iters/s | ms/iter | GDOF/s | efficiency |")
    print("|---|---|---|---|---|---|---|")
    for key in sorted(by):
        runs = by[key]
        n0 = min(runs)
        for n in sorted(runs):
            r = runs[n]
            if key[0] == "strong":
                eff = r["value"] / (runs[n0]["value"] * n / n0)
            else:
                eff = r["value"] / runs[n0]["value"]
            print(f"| {key[0]} | {'x'.join(map(str, key[1]))} | {n} | {r['value']:.1f} | {r['ms_per_step']:.4f} | "
                  f"{r.get('gdof_per_s', 0):.1f} | {eff:.2f} |")


if __name__ == "__main__":
    main()
