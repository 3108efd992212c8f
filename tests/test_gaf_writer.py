"""GAF writers (host formatter, rs-vgaligner_b200/host/gaf_writer.cpp) against
  (1) lines derived BY HAND from the reference's record code (src/align.rs:762-911, 913-930, 971-1027, 1096-1168) —
      independent of both writers,
  (2) the oracle's plain string writer on the oracle's result arrays (different code: no shared helpers), single- and
      multi-threaded.
CPU only: the result view is assembled from numpy arrays, no GPU call is made."""
import numpy as np
import pytest

import vgaligner_b200 as vg
from conftest import fixture_path
from oracle import binding as orc


class Arrays:
    """Result arrays under the C ABI's field names (include/vga.h, VgaResultsView)."""

    def __init__(self, n, **kw):
        z32, z64 = np.zeros(n, np.uint32), np.zeros(n + 1, np.uint64)
        self.n_anchors, self.chain_off, self.chain_anchor_off = z32.copy(), z64.copy(), np.zeros(1, np.uint64)
        self.ca_qb = self.ca_tb = self.ca_te = np.zeros(0, np.uint32)
        self.aln_status = np.zeros(n, np.uint8)
        self.aln_chain = self.aln_path_len = self.aln_start_off = self.aln_end_off = self.aln_n_aligned = z32.copy()
        self.aln_score, self.aln_cells = np.zeros(n, np.int32), np.zeros(n, np.uint64)
        self.aln_path_off, self.aln_run_off = z64.copy(), z64.copy()
        self.aln_path, self.aln_runs = np.zeros(0, np.uint64), np.zeros(0, np.uint32)
        for k, v in kw.items():
            setattr(self, k, np.asarray(v, dtype=getattr(self, k).dtype))


def text(ix, arrays, reads, which, threads=1):
    view = vg.view_from_arrays(arrays, reads.n)
    s, failed = vg.gaf_text(ix, view, which, reads.seqs, reads.offsets, reads.names, threads=threads)
    return s.decode(), failed


# test/test.gfa in id order: 1 CAAATAAG [0,8) 2 A [8] 3 G [9] 4 T [10] 5 C [11] 6 TTG [12,15) 7 A [15] 8 G [16]
# 9 AAAT [17,21) 10 AA [21,23) 11 TTTCT [23,28) 12 GG [28,30) 13 AGTTCTAT [30,38) ...
def test_chain_record_by_hand():
    """test.gfa, k=4, read seq0: one chain of anchors (qb, tb, te) = (1,15,20) (10,21,25) (11,22,26).
    from_chain (src/align.rs:892-910): query_start = first qb = 1, query_end = last qb + k = 15; per anchor
    "(>node:offset,>node:offset)," of the start and INCLUSIVE end position: 15 -> node 7 offset 0, 19 -> node 9 offset 2;
    21 -> node 10 offset 0, 24 -> node 11 offset 1; 22 -> node 10 offset 1, 25 -> node 11 offset 2.  mapq min(f64::MIN as
    u64, 254) = 0."""
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    reads = vg.read_seqs_from_file(fixture_path("single-read-test.fa"))
    a = Arrays(1, chain_off=[0, 1], chain_anchor_off=[0, 3], ca_qb=[1, 10, 11], ca_tb=[15, 21, 22], ca_te=[20, 25, 26])
    want = "seq0\t31\t1\t15\t+\t(>7:0,>9:2),(>10:0,>11:1),(>10:1,>11:2),\t0\t0\t0\t0\t0\t0\tta:Z:chain,n_anchors: 3\n"
    assert text(ix, a, reads, 0)[0] == want
    # the oracle's own pipeline produces exactly this chain, and its writer the same line
    oi = orc.OracleIndex(ix)
    o = oi.map_reads(reads)
    assert o.gaf(oi, reads, 0) == want
    assert text(ix, o, reads, 0)[0] == want


def test_alignment_record_by_hand():
    """generate_alignment (src/align.rs:1145-1167) on hand-made result arrays: read seq0 (31 bp) on the path
    >9>10>11 = AAAT|AA|TTTCT starting at offset 1 of node 9 ('AATAATTTCT'): 3 '=' (AAT vs read AAA? no: op codes are
    given, the writer does not re-derive them) — runs: 2 '=', 1 'X', 2 'I', 3 '=', 1 'D', 2 '='.
    cs: ':2' '*' + lower(graph base) + lower(read base) '+' inserted read bases ':3' '-' deleted graph base ':2';
    cigar: M runs merge '=' and 'X': 3M 2I 3M 1D 2M."""
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    reads = vg.read_seqs_from_file(fixture_path("single-read-test.fa"))
    # read = AAAAACGTTA...; graph walk from node 9 offset 1: A A T | A A | T T T C T
    # '='x2 consumes graph AA / read AA; 'X' graph T vs read A -> "*ta"; 'I'x2 read "AA" -> "+aa";
    # '='x3 graph AAT / read CGT (op codes are taken as given) -> ":3"; 'D' graph T -> "-t"; '='x2 -> ":2"
    runs = [(0 << 28) | 2, (1 << 28) | 1, (3 << 28) | 2, (0 << 28) | 3, (2 << 28) | 1, (0 << 28) | 2]
    a = Arrays(1, aln_status=[vg.ALN_OK], aln_path_off=[0, 3], aln_path=[9 * 2, 10 * 2, 11 * 2], aln_run_off=[0, 6],
               aln_runs=runs, aln_path_len=[9], aln_start_off=[1], aln_end_off=[3], aln_n_aligned=[8])
    want = "seq0\t31\t0\t31\t+\t>9>10>11\t9\t1\t3\t0\t8\t255\tas:i:-30 cs:Z::2*ta+aa:3-t:2,cg:Z:3M2I3M1D2M\n"
    got, failed = text(ix, a, reads, 1)
    assert got == want and failed == 0
    # a reverse handle prints '<id' and walks the reverse-complement label (seq_from_handle, src/index.rs:503-533):
    # node 11 reversed = AGAAA
    a = Arrays(1, aln_status=[vg.ALN_OK], aln_path_off=[0, 1], aln_path=[11 * 2 + 1], aln_run_off=[0, 2],
               aln_runs=[(2 << 28) | 2, (0 << 28) | 3], aln_path_len=[5], aln_start_off=[0], aln_end_off=[4], aln_n_aligned=[3])
    want = "seq0\t31\t0\t31\t+\t<11\t5\t0\t4\t0\t3\t255\tas:i:-30 cs:Z:-ag:3,cg:Z:2D3M\n"
    assert text(ix, a, reads, 1)[0] == want


def test_placeholder_and_failed_status():
    """from_placeholder_chain (src/align.rs:913-930; the reference's own expected string is at src/align.rs:1203-1231).
    A device alignment that did not complete is not a record type of the reference: placeholder line + counted."""
    ix = vg.Index.build(fixture_path("test.gfa"), 11)
    reads = vg.read_seqs_from_file(fixture_path("multiple-read-test.fa"))
    a = Arrays(reads.n)
    want = "".join("%s\t%d\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n" % (reads.names[i], len(reads.seq(i))) for i in range(reads.n))
    assert text(ix, a, reads, 0)[0] == want
    assert text(ix, a, reads, 1) == (want, 0)
    a.aln_status[:] = vg.ALN_TOO_WIDE
    assert text(ix, a, reads, 1) == (want, reads.n)
    a.aln_status[0] = vg.ALN_BAND_ERR
    assert text(ix, a, reads, 1) == (want, reads.n)


@pytest.mark.parametrize("graph,k,n,length", [("test.gfa", 4, 0, 0), ("graphs/19-MICB-4227-smooth.gfa", 11, 1500, 150),
                                              ("graphs/2-DRB1-3123.gfa", 11, 800, 150)])
def test_matches_the_oracle_writer(graph, k, n, length):
    ix = vg.Index.build(fixture_path(graph), k)
    reads = vg.simulate_reads(ix, n, length, 0.02, 0.002, 0.002, 0.1, 5) if n else vg.read_seqs_from_file(fixture_path("multiple-read-test.fa"))
    oi = orc.OracleIndex(ix)
    o = oi.map_reads(reads, n_threads=8)
    for which in (0, 1):
        want = o.gaf(oi, reads, which)
        assert text(ix, o, reads, which, threads=1) == (want, 0)
        assert text(ix, o, reads, which, threads=5) == (want, 0)       # ranges cut by output volume, concatenated in order


def test_generated_names():
    """Simulated read sets of millions of reads carry no name list: names are "read<base+i>"."""
    ix = vg.Index.build(fixture_path("graphs/19-MICB-4227-smooth.gfa"), 11)
    reads = vg.simulate_reads_mt(ix, 300, 150, seed=9, first_id=1000)
    oi = orc.OracleIndex(ix)
    o = oi.map_reads(reads, n_threads=4)
    view = vg.view_from_arrays(o, reads.n)
    got, _ = vg.gaf_text(ix, view, 1, reads.seqs, reads.offsets, None, name_base=1000)
    names = ["read%d" % (1000 + i) for i in range(reads.n)]

    class R:
        pass
    r = R()
    r.seqs, r.offsets, r.names, r.n = reads.seqs, reads.offsets, names, reads.n
    assert got.decode() == o.gaf(oi, r, 1)


def test_hand_derived_lines_of_the_worked_example():
    """test.gfa, k = 4, read seq0.  Both GAF writers (the host's parallel one and the oracle's) against lines
    derived BY HAND from the reference's record code, not from either writer:
      chains record      src/align.rs:892-910 — query_start = first anchor's query_begin (1), query_end = last anchor's query_end
                         (11 + 4), one "(>node:offset,>node:offset)," per anchor from AnchorPosOnGraph (src/chain.rs:89-127: the
                         chain's anchors are (qb,tb,te) = (1,15,20), (10,21,25), (11,22,26); linearisation starts of nodes 7, 9, 10,
                         11 are 15, 17, 21, 23, so 15 -> >7:0, 20-1 -> >9:2, 21 -> >10:0, 25-1 -> >11:1, 22 -> >10:1, 26-1 -> >11:2),
                         six zero columns (path_length .. mapping_quality = min(f64::MIN as u64, 254) = 0), notes
      alignments record  src/align.rs:1145-1167 — query_length = 31, 0, 31, '+', ">id" per path node, |abpoa_nodes|,
                         aln_start_offset, aln_end_offset, 0, n_aligned_bases, 255, "as:i:-30 " + cs + ",cg:Z:" + cigar
    The numbers of the alignments record are read from the result arrays; its cigar is rebuilt here from the run-length ops."""
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    reads = vg.read_seqs_from_file(fixture_path("single-read-test.fa"))
    oi = orc.OracleIndex(ix)
    o = oi.map_reads(reads, orc.default_params(), n_threads=1)
    chains_line = "seq0\t31\t1\t15\t+\t(>7:0,>9:2),(>10:0,>11:1),(>10:1,>11:2),\t0\t0\t0\t0\t0\t0\tta:Z:chain,n_anchors: 3\n"
    view = vg.view_from_arrays(o, reads.n)
    host_chains, _ = vg.gaf_text(ix, view, 0, reads.seqs, reads.offsets, reads.names)
    assert host_chains.decode() == chains_line
    assert o.gaf(oi, reads, 0) == chains_line
    # alignments record, assembled from the arrays by the rules above
    path = "".join((">" if h % 2 == 0 else "<") + str(h // 2) for h in o.aln_path.tolist())
    cigar, last = [], None
    for r in o.aln_runs.tolist():
        op, n = "MMDI"[r >> 28], r & 0xFFFFFFF
        if cigar and cigar[-1][0] == op:
            cigar[-1][1] += n
        else:
            cigar.append([op, n])
    cigar = "".join("%d%s" % (n, op) for op, n in cigar)
    host_aln, _ = vg.gaf_text(ix, view, 1, reads.seqs, reads.offsets, reads.names)
    for text in (host_aln.decode(), o.gaf(oi, reads, 1)):
        f = text.rstrip("\n").split("\t")
        assert f[:5] == ["seq0", "31", "0", "31", "+"] and f[5] == path == ">6>7>9>10>11>12>13>14>16>17"
        assert f[6:12] == [str(int(o.aln_path_len[0])), str(int(o.aln_start_off[0])), str(int(o.aln_end_off[0])), "0",
                           str(int(o.aln_n_aligned[0])), "255"]
        assert f[12].startswith("as:i:-30 ") and f[12].endswith(",cg:Z:" + cigar) and cigar == "5M3I6M1D17M"
        assert text.endswith("\n") and text.count("\n") == 1


def test_cs_prefix_switch_in_both_writers():
    """which | 2: the alignments record's cs string without its "cs:Z:" tag (rs-abPOA's exact format is [RECALLED]; the switch
    exists in the oracle's writer and in the host's so a comparison with the real binary flips a constant)."""
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    reads = vg.read_seqs_from_file(fixture_path("single-read-test.fa"))
    oi = orc.OracleIndex(ix)
    o = oi.map_reads(reads, orc.default_params(), n_threads=1)
    view = vg.view_from_arrays(o, reads.n)
    with_tag = vg.gaf_text(ix, view, 1, reads.seqs, reads.offsets, reads.names)[0].decode()
    without = vg.gaf_text(ix, view, 3, reads.seqs, reads.offsets, reads.names)[0].decode()
    assert "as:i:-30 cs:Z:" in with_tag and with_tag.replace("as:i:-30 cs:Z:", "as:i:-30 ") == without
    assert o.gaf(oi, reads, 3) == without and o.gaf(oi, reads, 1) == with_tag
