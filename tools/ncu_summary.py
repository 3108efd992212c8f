#!/usr/bin/env python
"""Summaries of ncu output for profiles/ (the raw .ncu-rep files stay in gpurun_out/, which is scratch).

    python tools/ncu_summary.py launches gpurun_out/launches_c2.csv          > profiles/rNN_launches_<wl>.txt
    python tools/ncu_summary.py full gpurun_out/poa_fill_full.ncu-rep        > profiles/rNN_full_<kernel>.txt

`launches` aggregates the `--metrics gpu__time_duration.sum` CSV per kernel (count, total, share);
`full` prints the counters DESIGN.md and bench.py quote (dram bytes, pipe utilisation, issue rate,
occupancy, stall reasons) for every launch captured with `--set full`.
"""
import collections
import csv
import subprocess
import sys

FULL_KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_bytes.sum", "l1tex__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
    "smsp__inst_executed_op_global_ld.sum", "smsp__inst_executed_op_global_st.sum",
]


def launches(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, mi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    unit = None
    for r in rows[1:]:
        try:
            v = float(r[mi].replace(",", ""))
        except ValueError:
            continue
        unit = r[ui]
        a = agg.setdefault(r[ki].split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += v
    scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit or "ns", 1e-6)
    tot = sum(a[1] for a in agg.values())
    print("# %s: %d launches, %.3f ms total (gpu__time_duration.sum, cold-cache serialised replay)" %
          (path, sum(a[0] for a in agg.values()), tot * scale))
    print("%-100s %6s %12s %7s" % ("kernel", "count", "ms", "share"))
    for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-100s %6d %12.4f %6.1f%%" % (k[:100], n, t * scale, 100.0 * t / tot if tot else 0.0))


def full(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        print("== %s  (id %s)" % (r[col["Kernel Name"]], r[col["ID"]]))
        for k in FULL_KEYS:
            if k in col:
                print("  %-70s %18s %s" % (k, r[col[k]], units[col[k]]))
        stalls = []
        for h, i in col.items():
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    stalls.append((float(r[i]), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        print("  warp stall reasons (warps per issue-active cycle):")
        for v, n in sorted(stalls, reverse=True)[:8]:
            print("    %-40s %8.3f" % (n, v))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
