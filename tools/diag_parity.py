#!/usr/bin/env python
"""GPU box diagnostic: CUDA path vs oracle on the first N reads of a bench workload; prints the reads whose alignment differs.
usage: python tools/diag_parity.py C5 821 [first_id]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import vgaligner_b200 as vg
from oracle import binding as orc

name, n = sys.argv[1], int(sys.argv[2])
first = int(sys.argv[3]) if len(sys.argv) > 3 else 0
wl = bench.WORKLOADS[name]
ix, _ = bench.build_index(vg, wl)
reads = vg.simulate_reads_mt(ix, n, wl["length"], *wl["err"], 0.1, wl["seed"], first_id=first)
g = vg.map_reads(ix, reads)
o = orc.OracleIndex(ix).map_reads(reads, orc.default_params(), n_threads=os.cpu_count())
bad = np.flatnonzero((g.aln_status != o.aln_status) | (g.aln_score != o.aln_score) | (g.aln_cells != o.aln_cells) |
                     (g.aln_path_len != o.aln_path_len) | (g.aln_n_aligned != o.aln_n_aligned))
print("reads", n, "differing", bad.size)
for r in bad[:40]:
    print("read", r, "len", int(reads.offsets[r + 1] - reads.offsets[r]), "status g/o", g.aln_status[r], o.aln_status[r], "score", g.aln_score[r], o.aln_score[r],
          "cells", g.aln_cells[r], o.aln_cells[r], "path_len", g.aln_path_len[r], o.aln_path_len[r], "n_anchors", g.n_anchors[r])
for f in bench.PARITY_FIELDS:
    a, b = getattr(g, f), getattr(o, f)
    print(f, "equal" if np.array_equal(a, b) else "DIFF shape %s %s" % (a.shape, b.shape))
