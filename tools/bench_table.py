#!/usr/bin/env python
"""Rows of DESIGN.md §7's table from the bench lines of one round: python tools/bench_table.py profiles/r01f"""
import json, sys
pre = sys.argv[1]
for W in ["C2", "C3a", "C3b", "C4", "C5"]:
    d = json.loads(open(f"{pre}_bench_{W}.json").read())
    r, l, c = d["roofline"], d.get("roofline_lookup", {}), d.get("cpu_baseline", {})
    print(f"| {W} | {d['value']:.4g} | {d['e2e']['value']:.4g} | {d['ms_per_step']:.1f} | {r['achieved']:.0f} ({100*r['frac']:.1f} %) | "
          f"{l.get('achieved', 0):.0f} ({100*l.get('frac', 0):.1f} %) | {c.get('value', 0):.3g} ({c.get('cores')} thr) |")
    print("    stage_ms", {k: round(v, 2) for k, v in d["stage_ms"].items()}, "steps", d["ms_steps"])
