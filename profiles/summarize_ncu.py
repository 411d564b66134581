#!/usr/bin/env python3
"""Turns ncu output into the small text summaries committed under profiles/.

  python profiles/summarize_ncu.py launches gpurun_out/launches.csv           > profiles/rNN_launches.txt
  python profiles/summarize_ncu.py full gpurun_out/k_render_full.ncu-rep      > profiles/rNN_k_render_full.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__cycles_elapsed.max", "smsp__cycles_active.avg", "smsp__inst_executed.sum", "sm__inst_issued.avg.per_cycle_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__grid_size", "launch__block_size",
    "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__maximum_warps_per_active_cycle_pct", "sass__inst_executed_shared_loads", "sass__inst_executed_shared_stores",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    k_name, k_val, k_grid, k_block = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        name = r[k_name].split("(")[0]
        a = agg.setdefault(name, {"n": 0, "ns": 0.0, "grid": r[k_grid], "block": r[k_block]})
        a["n"] += 1
        a["ns"] += float(r[k_val].replace(",", ""))
    total = sum(a["ns"] for a in agg.values())
    print(f"# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised per-launch times); {len(rows) - 1} launches")
    print(f"{'kernel':60s} {'launches':>8s} {'avg us':>10s} {'total us':>10s} {'share':>7s}  grid / block")
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
        print(f"{name[:60]:60s} {a['n']:8d} {a['ns'] / a['n'] / 1e3:10.2f} {a['ns'] / 1e3:10.1f} {100 * a['ns'] / total:6.1f}%  {a['grid']} / {a['block']}")


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(hdr, zip(units, vals)))
        print("# kernel:", d.get("Kernel Name", ("", "?"))[1], " grid", d.get("Grid Size", ("", "?"))[1], " block", d.get("Block Size", ("", "?"))[1])
        for k in KEYS:
            if k in d:
                print(f"{k:75s} {d[k][1]:>16s} {d[k][0]}")
        print("# warp issue-stall breakdown (average stalled warps per issue-active cycle), top 8")
        st = [(float(v[1].replace(",", "")), k) for k, v in d.items() if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio")]
        for v, k in sorted(st, reverse=True)[:8]:
            print(f"{k:75s} {v:16.3f}")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
