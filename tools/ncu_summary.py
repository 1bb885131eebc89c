#!/usr/bin/env python
"""Prints the handful of ncu raw-page metrics the design notes quote, per kernel of a report, and (with --traffic)
writes the per-launch DRAM traffic of every kernel into the JSON bench.py reads for roofline.traffic.
    python tools/ncu_summary.py <report.ncu-rep> [--traffic profiles/traffic.json --source profiles/<summary>.txt]
`--traffic` averages dram__bytes_read.sum + dram__bytes_write.sum over the captured launches of each kernel and keys
them by the kernel's base name (namespace and template arguments stripped)."""
import argparse
import csv
import json
import re
import subprocess
import sys

WANT = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__waves_per_multiprocessor", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "lts__t_sectors_op_atom.sum", "lts__t_sectors_op_red.sum",
        "smsp__average_warp_latency_per_inst_issued.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio"]

UNIT_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def base_name(kernel: str) -> str:
    k = re.sub(r"^void\s+", "", kernel)
    k = k.split("(")[0].split("<")[0]
    return k.split("::")[-1]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--traffic", help="JSON file to write the per-kernel DRAM bytes per launch into")
    ap.add_argument("--source", help="what to record as the provenance of --traffic (the committed summary's path)")
    ap.add_argument("--command", default="", help="the profiled command, recorded in --traffic")
    a = ap.parse_args()
    raw = subprocess.run(["ncu", "-i", a.report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    traffic = {}
    for r in rows[2:]:
        print("----")
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print(f"{w:90s} {r[i]} {units[i]}")
        name = base_name(r[hdr.index("Kernel Name")])
        tot = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = hdr.index(m)
            tot += float(r[i].replace(",", "")) * UNIT_BYTES[units[i]]
        traffic.setdefault(name, []).append(tot)
    if a.traffic:
        out = {"_doc": "dram__bytes_read.sum + dram__bytes_write.sum per launch (mean over the captured launches of each "
                       "kernel) from one `ncu --set full --clock-control none` capture; written by tools/ncu_summary.py, "
                       "read by bench.py for roofline.traffic",
               "source": a.source or a.report, "command": a.command,
               "launches_captured": {k: len(v) for k, v in traffic.items()}}
        out.update({k: int(round(sum(v) / len(v))) for k, v in traffic.items()})
        with open(a.traffic, "w") as f:
            json.dump(out, f, indent=1)
            f.write("\n")
        print(f"wrote {a.traffic}", file=sys.stderr)


if __name__ == "__main__":
    main()
