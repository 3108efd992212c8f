#!/usr/bin/env python
"""Turns the files of one tools/gpu_evidence.sh visit (gpurun_out/<tag>_*) into the committed summaries under profiles/ and
regenerates profiles/traffic.json (DRAM bytes per launch of the dominant kernels, from the ncu --set full captures of the CURRENT
build, with the git SHA of the tree that was profiled).      usage: python tools/make_profiles.py <tag>"""
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
    if len(rows) < 3:
        return None
    return dict(zip(rows[0], rows[2]))


def num(x):
    return float(str(x).replace(",", ""))


for f in sorted(os.listdir(G)):
    if not f.startswith(tag + "_"):
        continue
    src = os.path.join(G, f)
    if f.endswith(".json") or f.endswith("_pytest.log") or f.endswith(".txt") and "timeline" in f or f.endswith("_sha.txt"):
        shutil.copy(src, os.path.join(P, f))
    elif f.endswith(".csv") and "launches" in f:
        shutil.copy(src, os.path.join(P, f))
        with open(os.path.join(P, f[:-4] + ".txt"), "w") as o:
            subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), "launches", src], stdout=o)
    elif f.endswith(".ncu-rep"):
        with open(os.path.join(P, f[:-8] + ".txt"), "w") as o:
            subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_summary.py"), "full", src], stdout=o)

sha = open(os.path.join(G, tag + "_sha.txt")).read().strip() if os.path.exists(os.path.join(G, tag + "_sha.txt")) else "unknown"
traffic = {"source": "ncu --set full captures of tools/gpu_evidence.sh %s, tree %s" % (tag, sha), "_comment":
           "dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the kernel on the named workload size, and the algorithmic units of "
           "that same launch (cells / lookup bytes from a bench run with the same --reads): bench.py scales bytes per unit to its own launches"}
for wl, reads, key, rep, unit_key in (("C4", 250000, "poa_fill", "pair_C4", "poa_cells"), ("C2", 100000, "poa_fill", "pair_C2", "poa_cells"),
                                      ("C4", 1000000, "seed", "seed_count_C4", "lookup_bytes"),
                                      # C5: the fused long-read kernel is ONE launch per call; the call's cell count also holds the few
                                      # alignments that end on the wide kernel, so bytes per cell come out slightly low
                                      ("C5", 12500, "poa_fill", "long_C5", "poa_cells")):
    r = raw(os.path.join(G, "%s_full_%s.ncu-rep" % (tag, rep)))
    cj = os.path.join(G, "%s_counts_%s_%d.json" % (tag, wl, reads))
    if r is None or not os.path.exists(cj):
        continue
    units = json.load(open(cj))["counts"][unit_key]
    # ncu prints the byte counters with a unit column; --page raw gives plain numbers in the unit of the header row
    rd, wr = num(r["dram__bytes_read.sum"]), num(r["dram__bytes_write.sum"])
    unit_row = subprocess.run(["ncu", "-i", os.path.join(G, "%s_full_%s.ncu-rep" % (tag, rep)), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(l for l in unit_row.splitlines() if l.startswith('"')))
    units_of = dict(zip(rows[0], rows[1]))
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    rd *= scale.get(units_of["dram__bytes_read.sum"], 1.0)
    wr *= scale.get(units_of["dram__bytes_write.sum"], 1.0)
    if key == "seed":          # the stage is two kernels: seed_count (the probes) + seed_emit (positions read, anchors written)
        r2 = raw(os.path.join(G, "%s_full_seed_emit_C4.ncu-rep" % tag))
        if r2 is not None:
            out2 = subprocess.run(["ncu", "-i", os.path.join(G, "%s_full_seed_emit_C4.ncu-rep" % tag), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
            rows2 = list(csv.reader(l for l in out2.splitlines() if l.startswith('"')))
            u2 = dict(zip(rows2[0], rows2[1]))
            rd += num(r2["dram__bytes_read.sum"]) * scale.get(u2["dram__bytes_read.sum"], 1.0)
            wr += num(r2["dram__bytes_write.sum"]) * scale.get(u2["dram__bytes_write.sum"], 1.0)
    e = traffic.setdefault(wl, {})
    e[key + "_bytes_per_launch"] = rd + wr
    e[key + "_kernel"] = r.get("Kernel Name")
    e[key + "_units_in_launch"] = units
    e[key + "_note"] = "%s on %s with %d reads in one call: %.3f GB read + %.3f GB written; %s = %d" % (rep, wl, reads, rd / 1e9, wr / 1e9, unit_key, units)
    if wl == "C5":
        e[key + "_note"] += " (cells of the whole call: the fused kernel's launch plus the alignments that moved on to the wide kernel)"
json.dump(traffic, open(os.path.join(P, "traffic.json"), "w"), indent=1)
print(json.dumps(traffic, indent=1))
