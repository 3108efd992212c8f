"""Parity tests proper: the CUDA path through the C ABI (vga_map_batch) against the oracle on the
same seeded inputs, stage by stage.  Integer / index work is compared bit-exact; the f64 chain
scores are compared bit-exact as well (DESIGN.md explains why that is achievable).

Run on the B200 box: python -m pytest tests -m gpu
"""
import numpy as np
import pytest

import vgaligner_b200 as vg
from conftest import fixture_path, gfa_text
from oracle import binding as orc

pytestmark = pytest.mark.gpu


class FakeReads:
    """Reads assembled in Python (duck-types vgaligner_b200.Reads for both sides)."""

    def __init__(self, seqs, names=None):
        self.n = len(seqs)
        self.names = names or ["r%d" % i for i in range(len(seqs))]
        self.offsets = np.zeros(self.n + 1, dtype=np.uint64)
        if self.n:
            self.offsets[1:] = np.cumsum([len(s) for s in seqs])
        self.seqs = np.frombuffer("".join(seqs).encode(), dtype=np.uint8).copy() if self.n else np.zeros(0, np.uint8)
        self._h = None

    def seq(self, i):
        return self.seqs[int(self.offsets[i]):int(self.offsets[i + 1])].tobytes().decode()


def both(ix, reads, **kw):
    flags = vg.VGA_FLAG_KEEP_DEBUG | kw.pop("flags", 0)
    gp = vg.default_params(flags=flags, **kw)
    op = orc.default_params(flags=orc.FLAG_KEEP_DEBUG | (orc.FLAG_NO_ALIGN if flags & vg.VGA_FLAG_NO_ALIGN else 0), **kw)
    g = vg.map_reads(ix, reads, params=gp)
    o = orc.OracleIndex(ix).map_reads(reads, op, n_threads=8)
    return g, o


def assert_same(g, o, what):
    for name in what:
        a, b = getattr(g, name), getattr(o, name)
        assert a.shape == b.shape, "%s: shape %s vs %s" % (name, a.shape, b.shape)
        if not np.array_equal(a, b):
            bad = np.flatnonzero((a != b).reshape(a.shape[0], -1).any(axis=1))
            raise AssertionError("%s differs at %d entries, first %d: gpu=%s oracle=%s" %
                                 (name, bad.size, bad[0], a[bad[0]], b[bad[0]]))


SEED = ["n_anchors", "dbg_anchor_off", "dbg_qb", "dbg_tb", "dbg_te"]
SORT = ["dbg_sorted"]
CHAIN = ["dbg_f", "dbg_pred", "chain_off", "chain_anchor_off", "ca_id", "ca_qb", "ca_tb", "ca_te", "ca_score"]
SUB = ["dbg_sub_off", "dbg_sub_handles", "dbg_edge_off"]
ALN = ["aln_status", "aln_chain", "aln_score", "aln_cells", "aln_path_len", "aln_start_off", "aln_end_off", "aln_n_aligned",
       "aln_path_off", "aln_path", "aln_run_off", "aln_runs"]


def check_all(g, o):
    assert_same(g, o, SEED)
    assert_same(g, o, SORT)
    assert_same(g, o, CHAIN)
    assert_same(g, o, SUB)
    # edge lists: the oracle keeps find_nodes_edges_for_abpoa's emission order, the device a per-target
    # predecessor list; both orders are the same relation (compared as sorted pairs per read)
    for r in range(len(g.dbg_edge_off) - 1):
        a = g.dbg_edges[int(g.dbg_edge_off[r]):int(g.dbg_edge_off[r + 1])]
        b = o.dbg_edges[int(o.dbg_edge_off[r]):int(o.dbg_edge_off[r + 1])]
        assert sorted(map(tuple, a.tolist())) == sorted(map(tuple, b.tolist())), "edges of read %d" % r
    assert_same(g, o, ALN)


@pytest.mark.parametrize("k", [3, 4, 5, 6, 11])
@pytest.mark.parametrize("fa", ["single-read-test.fa", "multiple-read-test.fa"])
def test_c1_test_gfa(k, fa):
    """BASELINE config 1: test/test.gfa + the reference's FASTA fixtures, several k."""
    ix = vg.Index.build(fixture_path("test.gfa"), k)
    reads = vg.read_seqs_from_file(fixture_path(fa))
    g, o = both(ix, reads)
    check_all(g, o)
    oi = orc.OracleIndex(ix)
    assert g.gaf(0) == o.gaf(oi, reads, 0)
    assert g.gaf(1) == o.gaf(oi, reads, 1)
    if k == 11:
        assert g.gaf(1).split("\t")[2] == "*"       # placeholder record


def test_c1_whole_linearisation_chain():
    ix = vg.Index.build(fixture_path("test.gfa"), 11)
    reads = FakeReads([ix.seq_fwd, ix.seq_fwd[5:50]])
    g, o = both(ix, reads, min_anchors=2)
    check_all(g, o)
    assert g.ca_score[int(g.chain_anchor_off[1]) - 1] == 35.0


@pytest.mark.parametrize("graph,k,n,length", [
    ("graphs/2-DRB1-3123.gfa", 11, 3000, 150),
    ("graphs/19-MICB-4227-smooth.gfa", 11, 2000, 150),
    ("graphs/17-DRB1-3123-smooth.gfa", 11, 2000, 150),
])
def test_hla_graphs_short_reads(graph, k, n, length):
    """BASELINE configs 2 and 3 (scaled down): simulated 150 bp reads, 1% sub + 0.1% indel, 10% revcomp."""
    ix = vg.Index.build(fixture_path(graph), k)
    reads = vg.simulate_reads(ix, n, length, 0.01, 0.0005, 0.0005, 0.1, 77)
    g, o = both(ix, reads)
    check_all(g, o)
    oi = orc.OracleIndex(ix)
    assert g.gaf(0) == o.gaf(oi, reads, 0)
    assert g.gaf(1) == o.gaf(oi, reads, 1)
    assert (g.aln_status == vg.ALN_OK).mean() > 0.5


@pytest.mark.gpu
def test_small_trace_pool_waves_and_overflow(monkeypatch):
    """A trace pool far smaller than the batch needs: alignments run in several waves, the calibrated wave estimate
    overshoots at least sometimes and the pool-overflow path (IT_RETRY_POOL -> traceback, start over, rerun) is taken.
    Results must not depend on any of it."""
    ix = vg.Index.build(fixture_path("graphs/2-DRB1-3123.gfa"), 11)
    reads = vg.simulate_reads(ix, 3000, 150, 0.01, 0.0005, 0.0005, 0.1, 5)
    # 148 SMs x 64 x 8192 words are set aside for chunks in flight; leave ~12 M words (the batch needs > 100 M)
    monkeypatch.setenv("VGA_POOL_MAX_WORDS", str(148 * 64 * 8192 + 12_000_000))
    g, o = both(ix, reads)
    monkeypatch.delenv("VGA_POOL_MAX_WORDS")
    check_all(g, o)
    assert g.stats["poa_launches"] >= 3


@pytest.mark.gpu
def test_threaded_item_classification(monkeypatch):
    """Batches of >= 200 k alignments are classified on several host threads (per-thread histograms); forced here on a
    small batch, with a pool small enough that the class lists are cut into waves: results must be those of the oracle."""
    ix = vg.Index.build(fixture_path("graphs/2-DRB1-3123.gfa"), 11)
    reads = vg.simulate_reads(ix, 3000, 150, 0.01, 0.0005, 0.0005, 0.1, 11)
    monkeypatch.setenv("VGA_CLASSIFY_THREADS", "5")
    monkeypatch.setenv("VGA_POOL_MAX_WORDS", str(148 * 64 * 8192 + 40_000_000))
    g, o = both(ix, reads)
    monkeypatch.delenv("VGA_CLASSIFY_THREADS")
    monkeypatch.delenv("VGA_POOL_MAX_WORDS")
    check_all(g, o)


def test_cyclic_inverting_graph():
    """5-B3106 has back links and '-' orientations (reverse handles in the extension)."""
    ix = vg.Index.build(fixture_path("graphs/5-B3106.gfa"), 11)
    reads = vg.simulate_reads(ix, 500, 120, 0.01, 0.0005, 0.0005, 0.1, 3)
    g, o = both(ix, reads)
    check_all(g, o)


@pytest.fixture(scope="module")
def synth5m():
    """the 5 Mbp synthetic graph of configs C4 / C5, k=15 (same generator and seed as bench.py)"""
    return vg.Index.build_from_text(vg.synth_gfa(5000000, 100, 1), 15)


@pytest.fixture(scope="module")
def synth():
    return vg.Index.build_from_text(vg.synth_gfa(200000, 100, 1), 15)


def test_synthetic_short_reads(synth):
    """BASELINE config 4 (scaled down): synthetic bubbles, k=15, 150 bp reads."""
    reads = vg.simulate_reads(synth, 4000, 150, 0.01, 0.0005, 0.0005, 0.1, 2)
    g, o = both(synth, reads)
    check_all(g, o)
    assert (g.aln_status == vg.ALN_OK).mean() > 0.7


def test_synthetic_long_reads(synth):
    """BASELINE config 5 (scaled down): 10 kb reads at 10% error (wide-band POA)."""
    reads = vg.simulate_reads(synth, 24, 10000, 0.04, 0.03, 0.03, 0.1, 3)
    g, o = both(synth, reads)
    check_all(g, o)
    assert (g.aln_status == vg.ALN_OK).sum() >= 12


def test_long_reads_through_waves_instead_of_the_fused_kernel(synth, monkeypatch):
    """VGA_POA_FUSED=0: the shared-memory long-read kernel as fill waves + traceback launches (the path pool retries of the fused
    per-warp kernel fall back to), with the pool capped so that several waves are needed."""
    monkeypatch.setenv("VGA_POA_LONG", "1")
    monkeypatch.setenv("VGA_POA_FUSED", "0")
    reads = vg.simulate_reads(synth, 20, 10000, 0.04, 0.03, 0.03, 0.1, 17)
    g, o = both(synth, reads)
    check_all(g, o)
    monkeypatch.setenv("VGA_POOL_MAX_WORDS", str(148 * 64 * 8192 + 30_000_000))
    g2, _ = both(synth, reads)
    check_all(g2, o)
    assert g2.stats["poa_launches"] > g.stats["poa_launches"]


def test_fused_long_read_kernel_with_a_small_pool(synth, monkeypatch):
    """The fused fill + traceback kernel when arenas are small: alignments outgrow their arena into the shared overflow region,
    and when that is exhausted they come back as pool retries and take the wave path — same results."""
    monkeypatch.setenv("VGA_POA_LONG", "1")
    reads = vg.simulate_reads(synth, 40, 10000, 0.04, 0.03, 0.03, 0.1, 19)
    g, o = both(synth, reads)
    check_all(g, o)
    monkeypatch.setenv("VGA_POOL_MAX_WORDS", str(148 * 64 * 8192 + 60_000_000))
    g2, _ = both(synth, reads)
    check_all(g2, o)


@pytest.mark.parametrize("family", ["0", "1"])
def test_long_read_kernel_families(synth, monkeypatch, family):
    """Both fill kernels for alignments outside the short-read classes — register-resident SLIDE instances (0) and the
    shared-memory kernel with work proportional to the band (1) — on 10 kb, 3 kb and 1 kb reads (incl. several chains
    per read); the library picks one per batch by item count, the test forces each."""
    monkeypatch.setenv("VGA_POA_LONG", family)
    for n, length, err, kw in ((16, 10000, (0.04, 0.03, 0.03), {}), (60, 3000, (0.03, 0.02, 0.02), {}),
                               (200, 1000, (0.03, 0.01, 0.01), dict(best_n=3, min_anchors=2))):
        reads = vg.simulate_reads(synth, n, length, *err, 0.1, 11)
        g, o = both(synth, reads, **kw)
        check_all(g, o)
        assert (g.aln_status == vg.ALN_OK).sum() >= n // 3


def test_wide_band_kernel_on_every_shape(synth, monkeypatch):
    """The CTA-per-alignment kernel for bands beyond 1024 columns (k_poa_wide.cu), forced on alignments of every shape:
    150 bp reads on an HLA graph with merges, 1 kb reads with several chains, 10 kb reads."""
    monkeypatch.setenv("VGA_POA_WIDE", "1")
    ix = vg.Index.build(fixture_path("graphs/17-DRB1-3123-smooth.gfa"), 11)
    reads = vg.simulate_reads(ix, 600, 150, 0.01, 0.0005, 0.0005, 0.1, 5)
    g, o = both(ix, reads)
    check_all(g, o)
    for n, length, err, kw in ((12, 10000, (0.04, 0.03, 0.03), {}), (100, 1000, (0.03, 0.01, 0.01), dict(best_n=3, min_anchors=2))):
        reads = vg.simulate_reads(synth, n, length, *err, 0.1, 13)
        g, o = both(synth, reads, **kw)
        check_all(g, o)
        assert (g.aln_status == vg.ALN_OK).sum() >= n // 3


def test_bands_wider_than_1024_columns(synth5m):
    """Reverse-strand 10 kb reads on the 5 Mbp index still chain a few random k-mer hits; their global alignment has rows
    thousands of columns wide.  abPOA has no band limit, so neither has the CUDA path: the warp kernels hand such an
    alignment to the wide kernel (found by bench.py's parity sample on C5)."""
    reads = vg.simulate_reads_mt(synth5m, 160, 10000, 0.04, 0.03, 0.03, 0.1, 3)
    o = orc.OracleIndex(synth5m).map_reads(reads, orc.default_params(), n_threads=32)
    g = vg.map_reads(synth5m, reads)
    assert_same(g, o, FULL)
    wide = (o.aln_status == 1) & (o.aln_cells > 10000 * 600)
    assert wide.sum() >= 5 and (g.aln_status != vg.ALN_TOO_WIDE).all()


def test_compact_bucket_table_on_small_graphs(monkeypatch):
    """The 8-byte-entry bucket table (k <= 16; chosen by default only for tables beyond L2 size, i.e. the 5 Mbp index) forced on
    the small fixtures: inline single-position entries, runs in the position array, k-mers with up to 41 occurrences, and the
    16-byte slots forced on the same inputs — identical anchors either way."""
    for compact in ("1", "0"):
        monkeypatch.setenv("VGA_TABLE_COMPACT", compact)
        for graph, k in (("test.gfa", 4), ("graphs/2-DRB1-3123.gfa", 11), ("graphs/17-DRB1-3123-smooth.gfa", 11)):
            ix = vg.Index.build(fixture_path(graph), k)
            reads = vg.simulate_reads(ix, 800, 150 if k == 11 else 30, 0.01, 0.0005, 0.0005, 0.1, 21)
            g, o = both(ix, reads)
            assert_same(g, o, SEED + SORT + CHAIN)
        rep = vg.Index.build_from_text(gfa_text(["ACGT" * 40], []), 5)       # every 5-mer occurs ~40 times: runs longer than 30
        g, o = both(rep, FakeReads(["ACGTACGTACGTAC", "GTACGTACG"]))
        assert_same(g, o, SEED + SORT + CHAIN)


def test_range_checked_pair_class_rebases_its_planes(monkeypatch):
    """Class 1 (VGA_POA_CHK=1): short reads whose subgraph is thousands of rows long run on the 16x2 kernel with int16 planes
    re-centred on the row maximum (RowInfo.ref) instead of the 32-bit SLIDE kernel; same results."""
    monkeypatch.setenv("VGA_POA_CHK", "1")
    ix = vg.Index.build(fixture_path("graphs/17-DRB1-3123-smooth.gfa"), 11)
    reads = vg.simulate_reads(ix, 6000, 150, 0.01, 0.0005, 0.0005, 0.1, 23)
    g, o = both(ix, reads)
    check_all(g, o)
    assert ((o.aln_status == 1) & (o.aln_cells > 150 * 6000)).sum() >= 10      # subgraphs far beyond the a-priori int16 bound


@pytest.mark.parametrize("gran", ["VGA_FLAG_BAND_GRAN8", "VGA_FLAG_BAND_GRAN16"])
def test_band_granularity_switch(synth, gran):
    """The abPOA-compatibility switch for SIMD-vector band rounding, in every fill kernel family (16x2 short reads, long reads
    on the shared-memory kernel, HLA graph with merges): CUDA path = oracle at granularity 8 and 16."""
    vflag, oflag = getattr(vg, gran), getattr(orc, gran.replace("VGA_", ""))
    ix = vg.Index.build(fixture_path("graphs/17-DRB1-3123-smooth.gfa"), 11)
    cases = [(ix, vg.simulate_reads(ix, 1500, 150, 0.01, 0.0005, 0.0005, 0.1, 41), {}),
             (synth, vg.simulate_reads(synth, 1500, 150, 0.01, 0.0005, 0.0005, 0.1, 42), {}),
             (synth, vg.simulate_reads(synth, 10, 10000, 0.04, 0.03, 0.03, 0.1, 43), {}),
             (synth, vg.simulate_reads(synth, 80, 1000, 0.03, 0.01, 0.01, 0.1, 44), dict(best_n=3, min_anchors=2))]
    for index, reads, kw in cases:
        gp = vg.default_params(flags=vg.VGA_FLAG_KEEP_DEBUG | vflag, **kw)
        op = orc.default_params(flags=orc.FLAG_KEEP_DEBUG | oflag, **kw)
        g = vg.map_reads(index, reads, params=gp)
        o = orc.OracleIndex(index).map_reads(reads, op, n_threads=8)
        assert_same(g, o, ALN)
        base = orc.OracleIndex(index).map_reads(reads, orc.default_params(**kw), n_threads=8)
        assert int(o.aln_cells.sum()) > int(base.aln_cells.sum())          # the band really is wider


def test_medium_reads_and_best_n(synth):
    reads = vg.simulate_reads(synth, 300, 1000, 0.03, 0.01, 0.01, 0.0, 9)
    g, o = both(synth, reads, best_n=3, min_anchors=2)
    check_all(g, o)


def test_chains_only(synth):
    reads = vg.simulate_reads(synth, 500, 150, 0.01, 0.0005, 0.0005, 0.1, 4)
    g, o = both(synth, reads, flags=vg.VGA_FLAG_NO_ALIGN)
    assert_same(g, o, SEED + SORT + CHAIN + ["aln_status"])
    assert (g.aln_status == vg.ALN_PLACEHOLDER).all()


def test_edge_cases(synth):
    """Empty and ragged inputs, reads shorter than k, non-ACGT bytes, lowercase."""
    fwd = synth.seq_fwd
    seqs = ["", "A", fwd[100:114], fwd[100:115], fwd[200:350], "N" * 150, fwd[400:450] + "N" + fwd[451:550],
            fwd[600:750].lower(), fwd[800:950], "ACGT" * 40, fwd[1000:1400]]
    reads = FakeReads(seqs)
    g, o = both(synth, reads)
    check_all(g, o)
    assert g.n_anchors[0] == 0 and g.n_anchors[1] == 0 and g.n_anchors[2] == 0 and g.n_anchors[5] == 0
    assert g.n_anchors[7] == 0      # lowercase never matches (raw byte comparison, src/index.rs:314-321)


def test_empty_batch(synth):
    g = vg.map_reads(synth, FakeReads([]))
    assert g.n_anchors.size == 0 and g.chain_off.tolist() == [0]


def test_repeats_many_anchors():
    """Low-complexity graph: k-mers with many occurrences, anchors per read in the thousands."""
    unit = "ACGTTGCA"
    nodes = [unit * 40, "A", "C", unit * 40, "G", "T", unit * 30]
    edges = [(1, 2), (1, 3), (2, 4), (3, 4), (4, 5), (4, 6), (5, 7), (6, 7)]
    ix = vg.Index.build_from_text(gfa_text(nodes, edges), 7)
    reads = FakeReads([unit * 12, (unit * 20)[3:], unit * 3 + "A" + unit * 5])
    g, o = both(ix, reads)
    check_all(g, o)
    assert g.n_anchors.max() > 1000


def test_batch_independence(synth):
    """Results do not depend on how reads are batched (reads are independent units)."""
    reads = vg.simulate_reads(synth, 600, 150, 0.01, 0.0005, 0.0005, 0.1, 11)
    whole = vg.map_reads(synth, reads)
    half = FakeReads([reads.seq(i) for i in range(300, 600)], reads.names[300:])
    part = vg.map_reads(synth, half)
    assert np.array_equal(whole.aln_score[300:], part.aln_score)
    assert np.array_equal(whole.aln_status[300:], part.aln_status)
    b = int(whole.aln_run_off[300])
    assert np.array_equal(whole.aln_runs[b:], part.aln_runs)


def test_size_independent_properties(synth):
    """Properties that hold at any size: every aligned read's ops consume the whole read, path
    handles are valid node ids, runs alternate, scores are bounded by 2*len."""
    reads = vg.simulate_reads(synth, 20000, 150, 0.01, 0.0005, 0.0005, 0.1, 21)
    g = vg.map_reads(synth, reads)
    ok = g.aln_status == vg.ALN_OK
    assert ok.mean() > 0.7
    ops = g.aln_runs >> 28
    lens = (g.aln_runs & 0x0fffffff).astype(np.int64)
    read_of_run = np.repeat(np.arange(reads.n), np.diff(g.aln_run_off).astype(np.int64))
    consumed = np.bincount(read_of_run, weights=np.where(ops != 2, lens, 0), minlength=reads.n)
    assert np.array_equal(consumed[ok], np.full(ok.sum(), 150.0))
    graph_ops = np.bincount(read_of_run, weights=np.where(ops != 3, lens, 0), minlength=reads.n)
    assert np.array_equal(graph_ops[ok].astype(np.uint32), g.aln_path_len[ok])
    same_read = read_of_run[1:] == read_of_run[:-1]
    assert not (same_read & (ops[1:] == ops[:-1])).any()
    assert (g.aln_score[ok] <= 300).all()
    assert ((g.aln_path >> 1) >= 1).all() and ((g.aln_path >> 1) <= synth.view.n_nodes).all()
    st = g.stats
    assert st["kernel_launches"] > 0 and st["poa_cells"] >= int(g.aln_cells.sum())


def test_int32_peak_is_measurable():
    assert vg.measure_int32_peak() > 1e12


def test_c2_full_size_parity_and_repeatability():
    """BASELINE config 2 at its full size (100k x 150 bp on DRB1-3123): every alignment array bit-exact
    against the oracle, and a second call on the reused workspace returns the same bytes."""
    ix = vg.Index.build(fixture_path("graphs/2-DRB1-3123.gfa"), 11)
    reads = vg.simulate_reads(ix, 100000, 150, 0.01, 0.0005, 0.0005, 0.1, 77)
    g = vg.map_reads(ix, reads, chain_scores=True)
    o = orc.OracleIndex(ix).map_reads(reads, orc.default_params(), n_threads=32)
    assert_same(g, o, ["n_anchors", "chain_off", "chain_anchor_off", "ca_id", "ca_qb", "ca_tb", "ca_te", "ca_score"] + ALN)
    g2 = vg.map_reads(ix, reads, chain_scores=True)
    assert_same(g2, g, ["ca_id", "ca_score"] + ALN)
    g3 = vg.map_reads(ix, reads)           # default: no ids / scores cross the bus
    assert g3.ca_id.size == 0 and g3.ca_score.size == 0
    assert_same(g3, g, ["ca_qb", "ca_tb", "ca_te"] + ALN)
    assert (g.aln_status == vg.ALN_OK).mean() > 0.9


# ---------------------------------------------------------------------------------------------------------------
# BASELINE configs at their real sizes (the 5 Mbp index, whose table leaves L2, was never parity-checked)
# ---------------------------------------------------------------------------------------------------------------
FULL = ["n_anchors", "chain_off", "chain_anchor_off", "ca_qb", "ca_tb", "ca_te"] + ALN


def test_c4_real_index_short_reads(synth5m):
    """C4 on the real 5 Mbp index (≈10 M keys, table far larger than L2, 5 M pos2node entries): 50 k reads."""
    reads = vg.simulate_reads_mt(synth5m, 50000, 150, 0.01, 0.0005, 0.0005, 0.1, 2)
    g = vg.map_reads(synth5m, reads)
    o = orc.OracleIndex(synth5m).map_reads(reads, orc.default_params(), n_threads=32)
    assert_same(g, o, FULL)
    assert (g.aln_status == vg.ALN_OK).mean() > 0.8


def test_c5_real_index_long_reads_multi_wave(synth5m, monkeypatch):
    """C5 on the real index: 64 reads of 10 kb at 10 % error, with the trace pool capped so that the alignments run in
    several waves (the full-size run admits waves by the same rule against a 54 GB pool)."""
    reads = vg.simulate_reads_mt(synth5m, 64, 10000, 0.04, 0.03, 0.03, 0.1, 3)
    o = orc.OracleIndex(synth5m).map_reads(reads, orc.default_params(), n_threads=32)
    g = vg.map_reads(synth5m, reads)
    assert_same(g, o, FULL)
    monkeypatch.setenv("VGA_POOL_MAX_WORDS", str(148 * 64 * 8192 + 40_000_000))
    g2 = vg.map_reads(synth5m, reads)
    monkeypatch.delenv("VGA_POOL_MAX_WORDS")
    assert_same(g2, o, FULL)
    assert g2.stats["poa_launches"] > g.stats["poa_launches"]
    assert (g.aln_status == vg.ALN_OK).sum() >= 40


@pytest.mark.parametrize("graph,n", [("graphs/17-DRB1-3123-smooth.gfa", 125000), ("graphs/19-MICB-4227-smooth.gfa", 125000)])
def test_c3_full_shard_size(graph, n):
    """C3 at the per-GPU size of the 8-GPU configuration (1 M reads / 8): every array against the oracle."""
    ix = vg.Index.build(fixture_path(graph), 11)
    reads = vg.simulate_reads_mt(ix, n, 150, 0.01, 0.0005, 0.0005, 0.1, 77)
    g = vg.map_reads(ix, reads)
    o = orc.OracleIndex(ix).map_reads(reads, orc.default_params(), n_threads=32)
    assert_same(g, o, FULL)


@pytest.mark.parametrize("bandwidth,max_gap", [(1, 1000), (20, 1000), (32, 1000), (33, 1000), (50, 100), (50, 2000), (64, 1000), (65, 1000), (96, 500)])
def test_chaining_parameters(bandwidth, max_gap):
    """chain_anchors with other windows and gap limits than map_main's (bandwidth 50 is fixed there, -g sets max_gap): windows up to
    32 predecessors (near half only), 33..64 (far half evaluated on demand), above 64 (the generic f64 body); sorted order, DP
    scores, predecessors and chains against the oracle on reads with spurious anchors (k = 7 on DRB1-3123)."""
    ix = vg.Index.build(fixture_path("graphs/2-DRB1-3123.gfa"), 7)
    reads = vg.simulate_reads(ix, 300, 150, 0.02, 0.002, 0.002, 0.1, 31)
    g, o = both(ix, reads, bandwidth=bandwidth, max_gap=max_gap, flags=vg.VGA_FLAG_NO_ALIGN)
    assert_same(g, o, SEED + SORT + CHAIN)
    assert int(g.chain_off[-1]) > 100


@pytest.mark.parametrize("kw", [dict(match=1, mismatch=3, o1=5, e1=2, o2=20, e2=1), dict(wb=5, wf=0.03), dict(match=3, mismatch=5, wb=20, wf=0.0),
                                dict(match=2, mismatch=6, o1=6, e1=3, o2=30, e2=2, wb=8, wf=0.02)])
def test_other_scoring_and_band_parameters(synth, kw):
    """The POA with other scores, gap models and band parameters than abPOA's defaults (inside the range map_common admits): the
    16x2 kernel is compiled for the default gap model, so these go through the 32-lane group kernel (short reads) and the
    shared-memory / SLIDE kernels (long reads) — every alignment array against the oracle."""
    ix = vg.Index.build(fixture_path("graphs/2-DRB1-3123.gfa"), 11)
    g, o = both(ix, vg.simulate_reads(ix, 400, 150, 0.02, 0.002, 0.002, 0.1, 61), **kw)
    check_all(g, o)
    assert (g.aln_status == vg.ALN_OK).sum() > 300
    g, o = both(synth, vg.simulate_reads(synth, 24, 3000, 0.03, 0.02, 0.02, 0.1, 62), **kw)
    check_all(g, o)
    assert (g.aln_status == vg.ALN_OK).sum() >= 16


def test_failed_alignment_status_reaches_the_gaf_text(synth, monkeypatch):
    """A device alignment that did not complete (status other than OK / PLACEHOLDER; patched into real results here, the
    inputs of this suite never produce one) is written as a placeholder line and counted, never as a well-formed record."""
    reads = vg.simulate_reads(synth, 50, 150, 0.01, 0.0005, 0.0005, 0.0, 4)
    g = vg.map_reads(synth, reads)
    ok = np.flatnonzero(g.aln_status == vg.ALN_OK)
    assert ok.size > 10
    g.aln_status[ok[0]] = vg.ALN_TOO_WIDE
    view = vg.view_from_arrays(g, reads.n)
    text, failed = vg.gaf_text(synth, view, 1, reads.seqs, reads.offsets, reads.names)
    line = text.decode().split("\n")[ok[0]]
    assert failed == 1 and line.split("\t")[2:] == ["*"] * 9 + ["0", "*"]


def test_n_gpus_through_the_c_abi_equal_one_gpu(synth):
    """Multi-GPU behind the C ABI (vga_comm_unique_id + vga_index_broadcast, NCCL loaded at run time, no torch): the
    concatenated arrays and GAF texts of N ranks equal the single-GPU result.  Needs a box with at least 2 GPUs."""
    n_dev = vg.gpu_lib().vga_device_count()
    if n_dev < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    n_gpus = min(n_dev, 4)
    reads = vg.simulate_reads(synth, 6001, 150, 0.01, 0.0005, 0.0005, 0.1, 31)
    one = vg.map_reads(synth, reads)
    parts = vg.map_reads_multi(synth, reads, n_gpus)
    assert sum(p.view.n_reads for p in parts) == reads.n
    for f in ("n_anchors", "aln_status", "aln_score", "aln_cells", "aln_path_len", "aln_start_off", "aln_end_off", "aln_n_aligned",
              "ca_qb", "ca_tb", "ca_te", "aln_path", "aln_runs"):
        assert np.array_equal(np.concatenate([getattr(p, f) for p in parts]), getattr(one, f)), f
    for which in (0, 1):
        assert "".join(p.gaf(which) for p in parts) == one.gaf(which)


@pytest.mark.parametrize("graph,k", [("test.gfa", 3), ("test.gfa", 4), ("test.gfa", 11), ("graphs/19-MICB-4227-smooth.gfa", 11),
                                     ("graphs/2-DRB1-3123.gfa", 11), ("graphs/17-DRB1-3123-smooth.gfa", 11), ("synth", 15)])
def test_gpu_index_build_equals_the_host_build(graph, k, synth):
    """GPU index build (k-mer walks, stable sort + consecutive dedup, position table) against host/index_build.cpp — itself pinned
    on the reference's index vectors (tests/test_oracle_golden.py): every key, offset and table entry incl. the delimiters, on
    graphs with N runs (2-DRB1-3123), reverse links (17-DRB1-3123-smooth) and bubbles."""
    ix = synth if graph == "synth" else vg.Index.build(fixture_path(graph), k)
    keys, koff, ts, te, ms = vg.index_build_device(ix, k)
    n, nt = ix.view.n_keys, ix.view.n_table
    assert len(keys) == n and len(ts) == nt
    assert np.array_equal(keys, np.ctypeslib.as_array(ix.view.key_kmer2bit, shape=(n,)))
    hs, he, ho = ix.table()
    assert np.array_equal(koff, ho) and np.array_equal(ts, hs) and np.array_equal(te, he)
    assert ms > 0


def test_gpu_index_build_refuses_walk_explosions():
    """5-B3106 (cyclic, microsatellite loops with out-degree 17): 17^10 walks from one start — the device build returns
    VGA_ERR_UNSUPPORTED instead of running for minutes; such graphs are built on the host."""
    ix = vg.Index.build(fixture_path("graphs/5-B3106.gfa"), 11)
    with pytest.raises(vg.VgaError):
        vg.index_build_device(ix, 11)


def test_reverse_strand_anchors_and_chains():
    """`only_forward = false` (reference src/chain.rs:154-156): an index uploaded with every table position, anchors of both
    strands, Reverse-first sort, chaining within one orientation; chains GAF with '<' positions.  Half of the reads are reverse
    complemented, so most chains of those lie on the reverse strand; their alignment is refused (VGA_ALN_UNSUPPORTED), forward
    chains align as always."""
    for graph, k in (("graphs/19-MICB-4227-smooth.gfa", 11), ("graphs/17-DRB1-3123-smooth.gfa", 11), ("test.gfa", 4)):
        ix = vg.Index.build(fixture_path(graph), k)
        ix.upload_flags = vg.VGA_INDEX_ALL_ORIENTATIONS
        reads = vg.simulate_reads(ix, 1500, 150 if k == 11 else 30, 0.01, 0.0005, 0.0005, 0.5, 51)
        gp = vg.default_params(flags=vg.VGA_FLAG_KEEP_DEBUG | vg.VGA_FLAG_ALL_ORIENTATIONS)
        op = orc.default_params(flags=orc.FLAG_KEEP_DEBUG | orc.FLAG_ALL_ORIENTATIONS)
        g = vg.map_reads(ix, reads, params=gp)
        oi = orc.OracleIndex(ix)
        o = oi.map_reads(reads, op, n_threads=8)
        assert_same(g, o, SEED + SORT + CHAIN + ALN)
        assert g.gaf(0) == o.gaf(oi, reads, 0)
        if k == 11:
            assert (g.ca_tb >> 31).sum() > 1000 and (g.aln_status == vg.ALN_UNSUPPORTED).sum() > 100
            assert (g.aln_status == vg.ALN_OK).sum() > 100
            assert ",<" in g.gaf(0)
        # the forward-only default on the same reads: a flag / index mismatch is an error, not a silent fallback
        with pytest.raises(vg.VgaError):
            vg.map_reads(ix, reads, params=vg.default_params())
        ix.free_device()
