#!/usr/bin/env python
"""Executed warp-instructions per source line: joins the SASS page of an ncu report with nvdisasm line info.

    ncu -i X.ncu-rep --page source --csv --print-source sass --kernel-id ::regex:<kernel>:1 > sass.csv
    cuobjdump -xelf all build/<file>.o ; nvdisasm -g -c <file>.sm_100a.cubin > file.dis
    python tools/sass_lines.py sass.csv file.dis <mangled-kernel-substring> [source-file-substring]

Instructions are matched by position (the n-th SASS line of the kernel in both listings); each one is charged to
the innermost line of the named source file on its inline stack.
"""
import collections
import csv
import re
import sys


def main():
    sass_csv, dis, kern = sys.argv[1:4]
    srcsub = sys.argv[4] if len(sys.argv) > 4 else ".cu"
    rows = list(csv.reader(open(sass_csv)))
    hdr = next(i for i, r in enumerate(rows) if "Instructions Executed" in r)
    H = rows[hdr]
    ie, isrc = H.index("Instructions Executed"), H.index("Source")
    counts = [(r[isrc].strip(), int(r[ie])) for r in rows[hdr + 1:] if len(r) > ie and r[ie].isdigit()]
    lines = []
    inside, cur = False, None
    for ln in open(dis):
        if ln.startswith(".text."):
            inside = kern in ln
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)(.*)', ln)
        if m:
            if srcsub in m.group(1):
                cur = int(m.group(2))
            continue
        if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
            op = ln.split("*/", 1)[1].strip().rstrip(";").strip()
            lines.append((cur, op))
    if len(lines) != len(counts):
        print("# warning: %d SASS lines in the disassembly, %d in the report" % (len(lines), len(counts)))
    per = collections.Counter()
    ops = collections.Counter()
    tot = 0
    for (line, op), (_, n) in zip(lines, counts):
        per[line] += n
        ops[op.split()[1].split(".")[0] if op.startswith("@") else op.split()[0].split(".")[0]] += n
        tot += n
    print("# %d warp-instructions executed" % tot)
    src = None
    for ln in open(dis):
        m = re.match(r'\s*//## File "([^"]+)"', ln)
        if m and srcsub in m.group(1):
            src = m.group(1)
            break
    text = open(src).read().split("\n") if src else []
    acc = 0
    for line, n in sorted(per.items(), key=lambda x: -x[1])[:70]:
        acc += n
        t = text[line - 1].strip()[:100] if line and line <= len(text) else ""
        print("%6.2f%% %6.2f%%  L%-4s %s" % (100.0 * n / tot, 100.0 * acc / tot, line, t))
    print("# by opcode")
    for op, n in ops.most_common(25):
        print("%6.2f%%  %s" % (100.0 * n / tot, op))


if __name__ == "__main__":
    main()
