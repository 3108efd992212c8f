"""Debug helper: compare the GPU alignments with the oracle under each VGA_POA_DEBUG setting."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import vgaligner_b200 as vg
from oracle import binding as orc
gfa = sys.argv[1]; k = int(sys.argv[2]); n = int(sys.argv[3]); length = int(sys.argv[4])
ix = vg.Index.build(gfa, k)
reads = vg.simulate_reads(ix, n, length, 0.01, 0.0005, 0.0005, 0.1, 77)
o = orc.OracleIndex(ix).map_reads(reads, orc.default_params(), n_threads=8)
for dbg in sys.argv[5:]:
    os.environ["VGA_POA_DEBUG"] = dbg
    g = vg.map_reads(ix, reads)
    bad_st = np.flatnonzero(g.aln_status != o.aln_status)
    bad_sc = np.flatnonzero((g.aln_status == o.aln_status) & (g.aln_score != o.aln_score))
    bad_cells = np.flatnonzero(g.aln_cells != o.aln_cells)
    print("dbg", dbg, "status diff", bad_st.size, "score diff", bad_sc.size, "cells diff", bad_cells.size,
          "gpu status hist", np.bincount(g.aln_status, minlength=7).tolist(), "first", bad_st[:5].tolist(), bad_sc[:5].tolist())
    g.free()
