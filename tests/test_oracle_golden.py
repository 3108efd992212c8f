"""Pins the host index builder and the oracle against every known-answer vector the reference's own
tests hold for this path.  Each test cites the reference test it ports
(paths under the reference tree).  CPU only."""
import math
import os

import numpy as np
import pytest

import vgaligner_b200 as vg
from conftest import fixture_path, gfa_text
from oracle import binding as orc

F, R = "F", "R"


def build(nodes, edges, k):
    ix = vg.Index.build_from_text(gfa_text(nodes, edges), k)
    return ix, orc.OracleIndex(ix)


# create_simple_graph (src/index.rs:654-678, src/chain.rs:715-739)
G1 = (["A", "CT", "GA", "GCA"], [(1, 2), (1, 3), (2, 4), (3, 4)])
# the 7-node graph of test_wrong_index / multinode k-mers (src/index.rs:1490-1510, 1711-1726)
G7 = (["AAAAAAA", "TTT", "CCC", "GGGGGGG", "GGG", "CCC", "TTTTTTT"],
      [(1, 2), (1, 3), (2, 4), (3, 4), (4, 5), (4, 6), (5, 7), (6, 7)])


def h(i, rev=False):
    return 2 * i + (1 if rev else 0)


def test_forward_creation():
    """test_forward_creation (src/index.rs:760-824): linearisation, bit-vector, NodeRef triples."""
    ix, oi = build(*G1, 3)
    assert ix.seq_fwd == "ACTGAGCA"
    assert ix.seq_rev == "TGCTCAGT"
    assert ix.node_ref.tolist() == [[0, 0, 0], [1, 2, 1], [3, 4, 1], [5, 6, 2], [8, 8, 0]]
    bv = "".join("1" if any(oi.select(n) == p for n in range(1, 6)) else "0" for p in range(9))
    assert bv == "110101001"


def test_kmers_generation_counts():
    """test_kmers_generation (src/index.rs:826-840): 14 / 4 / 0 graph k-mers after dedup."""
    for k, expect in ((3, 14), (6, 4)):
        ix, _ = build(*G1, k)
        ts, _, _ = ix.table()
        assert int((ts != np.uint64(2 ** 64 - 1)).sum()) == expect
    ix, _ = build(*G1, 30)      # the reference uses k=100; the packed path caps k at 32, same outcome: no k-mer fits
    assert ix.n_kmers == 0


def test_simple_path_graph():
    """src/index.rs:842-890: ACG -> TTT -> CA."""
    ix, _ = build(["ACG", "TTT", "CA"], [(1, 2), (2, 3)], 3)
    assert ix.seq_fwd == "ACGTTTCA"
    assert ix.node_ref[1].tolist() == [3, 1, 1]
    assert ix.node_ref[2].tolist() == [6, 3, 1]
    ts, _, _ = ix.table()
    assert int((ts != np.uint64(2 ** 64 - 1)).sum()) == 12


def test_rank_and_inverse_rank():
    """test_inverse_rank (src/index.rs:1633-1650)."""
    _, oi = build(*G1, 3)
    assert [oi.rank(p) for p in range(8)] == [1, 2, 2, 3, 3, 4, 4, 4]
    assert [oi.inverse_rank(p) for p in range(8)] == [1, 1, 1, 2, 2, 3, 3, 4]


def test_node_id_and_handle_from_seqpos():
    """test_index_access_nodes (src/index.rs:1218-1243), test_handle_from_seqpos (:1424-1443)."""
    _, oi = build(*G1, 3)
    assert oi.handle_from_seqpos(False, 0) >> 1 == 1
    assert oi.handle_from_seqpos(False, 2) >> 1 == 2
    assert oi.handle_from_seqpos(True, 0) >> 1 == 4
    assert oi.handle_from_seqpos(False, 0) == h(1)
    assert oi.handle_from_seqpos(True, 0) == h(4, True)


def test_edges_and_reverse_handles():
    """src/index.rs:1260-1367: Index.edges per node, incoming/outgoing incl. reverse handles."""
    ix, oi = build(*G1, 3)
    e = ix.edges.tolist()
    nr = ix.node_ref
    per_node = [e[int(nr[i][1]):int(nr[i + 1][1])] for i in range(4)]
    assert per_node == [[h(2), h(3)], [h(1), h(4)], [h(1), h(4)], [h(2), h(3)]]
    assert oi.outgoing(h(1)) == [h(2), h(3)] and oi.incoming(h(1)) == []
    assert oi.incoming(h(4)) == [h(2), h(3)] and oi.outgoing(h(4)) == []
    assert oi.incoming(h(1, True)) == [h(3, True), h(2, True)]
    assert oi.outgoing(h(4, True)) == [h(3, True), h(2, True)]
    assert oi.outgoing(h(2, True)) == [h(1, True)] and oi.incoming(h(2, True)) == [h(4, True)]


def test_index_access_positions():
    """test_index_access / test_index_access_2 (src/index.rs:1108-1170)."""
    _, oi = build(*G1, 3)
    assert oi.positions("ACT") == [((F, 0), (F, 3))]
    _, oi2 = build(["TTT", "AAA"], [(1, 2)], 3)
    assert oi2.positions("TTT") == [((F, 0), (F, 3)), ((R, 0), (R, 3))]
    assert oi.positions("AC") == [] and oi.positions("NNN") == []


def test_multinode_kmers():
    """test_index_contains_multinode_kmers (src/index.rs:1668-1732)."""
    _, oi = build(*G1, 5)
    assert len(oi.positions("ACTGC")) > 0 and len(oi.positions("CTGCA")) > 0
    _, o2 = build(["ACG", "C", "G", "TTTTT"], [(1, 2), (1, 3), (2, 4), (3, 4)], 5)
    assert o2.positions("ACGGT")[0] == ((F, 0), (F, 6))
    assert o2.positions("GCTTT")[0] == ((F, 2), (F, 8))
    assert o2.positions("CTTTT")[0] == ((F, 3), (F, 9))
    _, o3 = build(["ACG", "C", "G", "TTTTT", "TA", "CG", "TTT"],
                  [(1, 2), (1, 3), (2, 4), (3, 4), (4, 5), (4, 6), (5, 7), (6, 7)], 5)
    assert o3.positions("TTCGT")[0] == ((F, 8), (F, 15))


def test_node_ids_on_both_strands():
    """test_wrong_index (src/index.rs:1490-1608): node_id(F, seq_idx_i) = i+1, node_id(R, seq_idx_i) = n-i."""
    ix, oi = build(*G7, 11)
    n = 7
    for i in range(n):
        p = int(ix.node_ref[i][0])
        assert oi.handle_from_seqpos(False, p) >> 1 == i + 1
        assert oi.handle_from_seqpos(True, p) >> 1 == n - i


def test_seq_from_handle_and_select():
    """src/index.rs:1394-1422, 1652-1666."""
    ix, oi = build(*G1, 3)
    rc = {"A": "T", "C": "G", "G": "C", "T": "A"}
    for i, s in enumerate(G1[0]):
        assert oi.seq_from_handle(h(i + 1)) == s
        assert oi.seq_from_handle(h(i + 1, True)) == "".join(rc[c] for c in reversed(s))
        assert oi.select(i + 1) == int(ix.node_ref[i][0])
    assert oi.select(5) == 8         # the end-marker bit


def test_anchors_found():
    """anchors_found / anchors_found_2 / test_no_anchors* (src/chain.rs:741-777, 903-918)."""
    _, oi = build(*G1, 3)
    assert len(oi.anchors("ACTGCA", True)) == 4
    assert len(oi.anchors("AGAGC", True)) == 3
    assert len(oi.anchors("ACTGCA", False)) >= 4
    assert oi.anchors("AAATTT", False) == []
    assert oi.anchors("", False) == []
    _, o2 = build(["AAAAAAAAAAA", "C", "G", "TTTTTTTTTTTT"], [(1, 2), (1, 3), (2, 4), (3, 4)], 11)
    assert len(o2.anchors("AAAAACTTTTTT", True)) == 2


def test_simple_anchor_coordinates():
    """test_simple_anchors (src/chain.rs:806-823)."""
    _, oi = build(["ACT"], [], 3)
    assert oi.anchors("ACT", False) == [(0, 3, (F, 0), (F, 3))]


def test_reverse_strand_anchors():
    """test_simple_anchors_reverse / _2 (src/chain.rs:826-888)."""
    g = (["AAA", "CCC", "GGG", "AAA"], [(1, 2), (1, 3), (2, 4), (3, 4)])
    _, oi = build(*g, 3)
    a = oi.anchors("TTT", False)
    assert len(a) == 2
    for anchor, node in zip(a, (4, 1)):
        hb = oi.handle_from_seqpos(anchor[2][0] == R, anchor[2][1])
        he = oi.handle_from_seqpos(anchor[3][0] == R, anchor[3][1] - 1)
        assert hb == he == h(node, True)
    _, o9 = build(*g, 9)
    a = o9.anchors("TTTCCCTTT", False)
    assert len(a) == 1
    assert o9.handle_from_seqpos(True, a[0][2][1]) == h(4, True)
    assert o9.handle_from_seqpos(True, a[0][3][1] - 1) == h(1, True)


def test_score_anchor_known_answer():
    """test_score_anchors (src/chain.rs:1000-1035): expects -f64::MAX."""
    a = (35, 46, 3907, 3918, 31.397)
    b = (49, 60, 3906, 3918, 0.0)
    assert orc.score_anchor(a, b, 11, 100) == -1.7976931348623157e308


def test_score_anchor_arithmetic():
    """gamma_c and rounding of score_anchor (src/chain.rs:313-363) against a direct evaluation."""
    a = (0, 15, 100, 115, 15.0)
    b = (20, 35, 127, 142, 0.0)
    g = 27 - 20
    expect = round((15.0 + 15 - (0.01 * 15 * g + 0.5 * math.log2(g))) * 1000.0) / 1000.0
    assert orc.score_anchor(a, b, 15, 1000) == expect
    assert orc.score_anchor(a, (20, 35, 2000, 2015, 0.0), 15, 1000) == -1.7976931348623157e308   # gap > max_gap


def test_chains_on_whole_linearisation():
    """test_chains_2 (src/chain.rs:945-976): test.gfa k=11, read = seq_fwd, min_anchors=2 -> a chain;
    SURVEY App. D adds: 15 anchors -> one 15-anchor chain with f = 35.0."""
    ix = vg.Index.build(fixture_path("test.gfa"), 11)
    oi = orc.OracleIndex(ix)
    seq = ix.seq_fwd

    class One:
        n = 1
        names = ["whole"]
        seqs = np.frombuffer(seq.encode(), dtype=np.uint8)
        offsets = np.array([0, len(seq)], dtype=np.uint64)

    r = oi.map_reads(One, orc.default_params(min_anchors=2, flags=orc.FLAG_NO_ALIGN))
    assert r.n_anchors[0] > 0 and r.chain_off[1] >= 1
    assert r.n_anchors[0] == 15
    assert int(r.chain_anchor_off[1]) == 15 and r.ca_score[-1] == 35.0


def test_worked_example_k4():
    """SURVEY App. D worked example on test.gfa, k=4, read seq0 (hand emulation of the reference)."""
    ix = vg.Index.build(fixture_path("test.gfa"), 4)
    oi = orc.OracleIndex(ix)
    assert ix.seq_fwd == "CAAATAAGAGTCTTGAGAAATAATTTCTGGAGTTCTATATATATATCCAACTCTCTG"
    assert ix.n_kmers == 92 and ix.n_kmer_pos - ix.n_kmers == 156
    reads = vg.read_seqs_from_file(fixture_path("single-read-test.fa"))
    r = oi.map_reads(reads, orc.default_params(flags=orc.FLAG_KEEP_DEBUG))
    got = list(zip(r.dbg_qb.tolist(), r.dbg_tb.tolist(), r.dbg_te.tolist()))
    assert got == [(0, 15, 20), (1, 15, 20), (9, 1, 5), (9, 17, 21), (10, 18, 24), (10, 21, 25), (11, 8, 14),
                   (11, 19, 25), (11, 22, 26), (12, 10, 15), (13, 12, 17), (25, 0, 4), (26, 15, 20), (27, 15, 20)]
    assert r.dbg_sorted.tolist() == [11, 2, 6, 9, 10, 0, 1, 12, 13, 3, 4, 5, 7, 8]
    assert r.dbg_f.tolist() == [4.0, 4.0, 4.639, 5.639, 6.599, 4.0, 4.0, 7.694, 7.578, 4.0, 5.468, 6.84, 6.468, 7.84]
    assert r.ca_id.tolist() == [1, 5, 8]
    assert (r.dbg_sub_handles >> 1).tolist() == list(range(6, 19))
    assert r.dbg_edges.tolist() == [[0, 1], [0, 2], [1, 3], [2, 3], [3, 4], [3, 5], [4, 5], [5, 6], [5, 7], [6, 7],
                                    [7, 8], [7, 9], [8, 10], [9, 10], [10, 11], [10, 12]]


@pytest.mark.parametrize("k,expect_chain", [(11, False), (7, False), (6, False), (5, False), (4, True)])
def test_placeholder_behaviour_on_c1(k, expect_chain):
    """BASELINE config 1 yields a placeholder for k >= 5."""
    ix = vg.Index.build(fixture_path("test.gfa"), k)
    oi = orc.OracleIndex(ix)
    reads = vg.read_seqs_from_file(fixture_path("single-read-test.fa"))
    r = oi.map_reads(reads)
    assert (r.chain_off[1] > 0) == expect_chain
    if not expect_chain:
        assert r.gaf(oi, reads, 0) == "seq0\t31\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n"
        assert r.gaf(oi, reads, 1) == "seq0\t31\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n"


def test_placeholder_gaf_string():
    """test_to_string_placeholder (src/align.rs:1203-1231)."""
    ix, oi = build(*G1, 3)

    class One:
        n = 1
        names = ["Read1"]
        seqs = np.frombuffer(b"AAACTA", dtype=np.uint8)
        offsets = np.array([0, 6], dtype=np.uint64)

    r = oi.map_reads(One, orc.default_params(min_anchors=50))
    assert r.gaf(oi, One, 1) == "Read1\t6\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n"


def test_literal_extension_equals_memoised():
    """extend_range_chain_2 as written (no visited set) and the memoised walk return the same handles."""
    ix = vg.Index.build(fixture_path("graphs/19-MICB-4227-smooth.gfa"), 11)
    oi = orc.OracleIndex(ix)
    reads = vg.simulate_reads(ix, 60, 60, 0.02, 0.002, 0.002, 0.0, 5)
    a = oi.map_reads(reads, orc.default_params(flags=orc.FLAG_KEEP_DEBUG))
    b = oi.map_reads(reads, orc.default_params(flags=orc.FLAG_KEEP_DEBUG | orc.FLAG_LITERAL_EXTEND))
    assert a.dbg_sub_handles.tolist() == b.dbg_sub_handles.tolist()
    assert a.dbg_edges.tolist() == b.dbg_edges.tolist()
    assert a.aln_runs.tolist() == b.aln_runs.tolist()
    assert (a.aln_status == 1).sum() > 10


def test_all_orientations_through_map_reads():
    """ORC_FLAG_ALL_ORIENTATIONS = anchors_for_query(.., only_forward = false) (src/chain.rs:154-156) inside the per-read driver:
    the reverse complement of a read that chains on the forward strand chains on the reverse strand, with the orientation in
    bit 31 of ca_tb / ca_te, and is not aligned; the chains record spells reverse handles with '<' (src/align.rs:856-888)."""
    ix = vg.Index.build(fixture_path("graphs/19-MICB-4227-smooth.gfa"), 11)
    fwd = ix.seq_fwd[1000:1150]
    rc = fwd[::-1].translate(str.maketrans("ACGT", "TGCA"))

    class R:
        n = 2
        names = ["f", "r"]
        seqs = np.frombuffer((fwd + rc).encode(), dtype=np.uint8).copy()
        offsets = np.array([0, 150, 300], dtype=np.uint64)
    oi = orc.OracleIndex(ix)
    a = oi.map_reads(R, orc.default_params(), n_threads=1)
    b = oi.map_reads(R, orc.default_params(flags=orc.FLAG_ALL_ORIENTATIONS), n_threads=1)
    assert a.chain_off.tolist() == [0, 1, 1] and a.aln_status.tolist() == [1, 0]          # forward only: the reverse read is a placeholder
    assert b.chain_off.tolist()[1] >= 1 and b.chain_off.tolist()[2] > b.chain_off.tolist()[1]
    lo = int(b.chain_anchor_off[b.chain_off[1]])
    assert (b.ca_tb[lo:] >> 31).all() and (b.ca_te[lo:] >> 31).all()
    assert b.aln_status.tolist() == [1, 5]                                                  # VGA_ALN_UNSUPPORTED = 5
    line = b.gaf(oi, R, 0).split("\n")[-2]
    assert line.startswith("r\t150\t") and "(<" in line and ",<" in line
