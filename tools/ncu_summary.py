#!/usr/bin/env python
"""Turn an .ncu-rep (brought back in gpurun_out/) into a short committed summary under profiles/.

    python tools/ncu_summary.py gpurun_out/prof_spmv.ncu-rep profiles/r1_spmv_simple.md "note..."
"""
import csv
import io
import json
import re
import subprocess
import sys
from pathlib import Path

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "lts__t_bytes.sum",
    "l1tex__t_bytes.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    # issue-boundness (profiles/ROOFLINE.md R2.6): instructions, issue-slot utilisation, fp64 pipe, top stall reasons
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg",
    "smsp__inst_executed_op_tma_ld.sum", "smsp__average_warp_latency_per_inst_issued.ratio",
]


STALL = re.compile(r"smsp__average_warps_issue_stalled_(\w+)_per_issue_active\.ratio$")


def main():
    rep, out = Path(sys.argv[1]), Path(sys.argv[2])
    note = sys.argv[3] if len(sys.argv) > 3 else ""
    txt = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    peaks = json.loads(Path("MEASURED_PEAKS.json").read_text()) if Path("MEASURED_PEAKS.json").exists() else {}
    lines = [f"# ncu summary: `{rep.name}`", "", note, "",
             "Captured with `ncu --set full --clock-control none --import-source on` on a B200 via gpurun "
             "(cold caches, serialised launches: compare shares, not absolutes).", ""]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        lines += [f"## `{d.get('Kernel Name', '?')}` (launch id {d.get('ID', '?')})", "", "| metric | value | unit |",
                  "|---|---|---|"]
        for w in WANT:
            if w in d:
                lines.append(f"| `{w}` | {d[w]} | {units[hdr.index(w)]} |")
        # warp states: average number of warps per scheduler in each stall state per issue cycle, largest first
        stalls = []
        for k in hdr:
            m = STALL.match(k)
            if m and d.get(k):
                try:
                    stalls.append((float(d[k].replace(",", "")), m.group(1)))
                except ValueError:
                    pass
        for v, name in sorted(stalls, reverse=True)[:6]:
            lines.append(f"| warps stalled: `{name}` (per issue-active cycle) | {v:.2f} | warps |")
        try:
            dur = float(d["gpu__time_duration.sum"].replace(",", ""))
            unit = units[hdr.index("gpu__time_duration.sum")]
            dur_s = dur * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(unit, 1e-9)

            def b(name):
                v = float(d[name].replace(",", ""))
                u = units[hdr.index(name)]
                return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            tot = b("dram__bytes_read.sum") + b("dram__bytes_write.sum")
            gbs = tot / dur_s / 1e9
            frac = f" = {gbs / peaks['hbm_gbs']:.2f} of measured copy bandwidth ({peaks['hbm_gbs']} GB/s)" if peaks else ""
            lines += ["", f"DRAM traffic {tot / 1e6:.1f} MB in {dur_s * 1e6:.1f} us -> {gbs:.0f} GB/s{frac}.", ""]
        except Exception as e:  # pragma: no cover
            lines += ["", f"(could not derive bandwidth: {e})", ""]
    out.write_text("\n".join(lines) + "\n")
    print(out)


if __name__ == "__main__":
    main()
