#!/usr/bin/env python
"""Prints the measurement table of DESIGN.md §7 / BASELINE.md from the bench lines of one evidence visit (profiles/<tag>_bench_*.json)."""
import json
import os
import sys

tag = sys.argv[1]
P = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles")
rows = []
d = json.load(open(os.path.join(P, tag + "_bench_default.json")))
rows.append(("C4 synthetic 5 Mbp, k=15, 10 M × 150 bp", d))
rows.append(("C5 synthetic 5 Mbp, 100 k × 10 kb @10 %", d["long_reads"]))
for w, name in (("C2", "C2 DRB1-3123, k=11, 100 k × 150 bp"), ("C3a", "C3a 17-DRB1-3123-smooth, 1 M × 150 bp"), ("C3b", "C3b 19-MICB-4227-smooth, 1 M × 150 bp")):
    f = os.path.join(P, "%s_bench_%s.json" % (tag, w))
    if os.path.exists(f):
        rows.append((name, json.load(open(f))))
print("| workload (1 GPU, total reads) | reads/s device-resident | reads/s e2e (pinned / pageable host buffers) | e2e + both GAF texts | ms/step | POA fill GCUPS (frac of INT32 roofline) | lookup GB/s (frac of HBM) | aligned | CPU port, %d threads (1 thread) | parity sample |" % d["cpu_baseline"]["cores"])
print("|---|---|---|---|---|---|---|---|---|---|")
fmt = lambda v: ("%.2f M" % (v / 1e6)) if v >= 1e6 else ("%.1f k" % (v / 1e3))
for name, r in rows:
    cb, ps = r.get("cpu_baseline", {}), r.get("parity_sample", {})
    print("| %s | **%s** | %s / %s | %s | %.1f | %.0f (%.1f %%) | %.0f (%.1f %%) | %.1f %% | %s (%s) | %d reads, %s |" % (
        name, fmt(r["value"]), fmt(r["e2e"]["value"]), fmt(r["e2e"]["pageable_value"]), fmt(r["e2e_gaf"]["value"]) if r.get("e2e_gaf") else "—",
        r["ms_per_step"], r["roofline_poa"]["achieved"], 100 * r["roofline_poa"]["frac"], r["roofline_lookup"]["achieved"],
        100 * r["roofline_lookup"]["frac"], 100 * r["aligned_fraction"], fmt(cb.get("value", 0)), fmt(cb.get("t1_value", 0)),
        ps.get("n", 0), "equal" if ps.get("equal") else "DIFFERENT"))
print()
print("| workload | h2d | seed | sort+chain | chain gather | subgraph | POA fill | traceback | d2h | total (ms per step) |")
print("|---|---|---|---|---|---|---|---|---|---|")
for name, r in rows:
    s = r["stage_ms"]
    print("| %s | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f |" % (name.split(",")[0].split(" ")[0], s["ms_h2d"], s["ms_seed"], s["ms_sort"], s["ms_chain"],
                                                                       s["ms_subgraph"], s["ms_poa"], s["ms_trace"], s["ms_d2h"], s["ms_total"]))
