#!/usr/bin/env python
"""Fills the measurement tables of DESIGN.md and BASELINE.md (between the TABLE markers) from profiles/<tag>_bench_*.json.
usage: python tools/fill_docs.py <tag>"""
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
table = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "design_table.py"), tag], capture_output=True, text=True, check=True).stdout.strip()
ref = json.load(open(os.path.join(ROOT, "profiles", tag + "_bench_ref.json")))
ref_s = "C4 %.1f k reads/s, C5 %.0f reads/s on %d host threads (%s)" % (ref["value"] / 1e3, ref["long_reads"]["value"], ref["cpu_baseline"]["cores"],
                                                                     ref["cpu_baseline"]["sample"])
block = "<!-- TABLE %s -->\n%s\n<!-- /TABLE -->" % (tag, table)
for name in ("DESIGN.md", "BASELINE.md"):
    p = os.path.join(ROOT, name)
    s = open(p).read()
    if "@@TABLE@@" in s:
        s = s.replace("@@TABLE@@", block)
    else:
        s = re.sub(r"<!-- TABLE [^>]* -->.*?<!-- /TABLE -->", lambda m: block, s, flags=re.S)
    s = s.replace("@@TAG@@", tag).replace("@@REF@@", ref_s)
    s = re.sub(r"(Results of the last evidence visit \(`profiles/)[^_`]+(_\*`)", r"\g<1>%s\g<2>" % tag, s)
    s = re.sub(r"(Reference arm alone \(`profiles/)[^_`]+(_bench_ref.json`\): ).*?\.\n", lambda m: m.group(1) + tag + m.group(2) + ref_s + ".\n", s)
    open(p, "w").write(s)
print(table)
