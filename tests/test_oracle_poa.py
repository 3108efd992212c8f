"""Self-consistency of the oracle's POA definition (the part of the path whose parity is UNPINNED:
the reference delegates it to ab_poa / abPOA, reference src/align.rs:190-202, and pins no result).

An independent, unbanded pure-Python DP checks the recurrences; the traceback is checked by
re-scoring its own operations."""
import random

import numpy as np
import pytest

from oracle import binding as orc

M, X, D, I = 0, 1, 2, 3


def runs_of(r, i=0):
    b, e = int(r.aln_run_off[i]), int(r.aln_run_off[i + 1])
    return [(int(x) >> 28, int(x) & 0x0fffffff) for x in r.aln_runs[b:e]]


def gap(n, P):
    return min(P.o1 + n * P.e1, P.o2 + n * P.e2)


def rescore(runs, P):
    s = 0
    for op, n in runs:
        s += n * P.match if op == M else -n * P.mismatch if op == X else -gap(n, P)
    return s


def brute_force(labels, edges, q, P):
    """Unbanded global sequence-to-graph DP with the convex gap min(o1+g*e1, o2+g*e2): best score
    over paths from a predecessor-less node's first base to a successor-less node's last base."""
    NEG = -10 ** 9
    rows, preds = [], []
    first_row, last_row = {}, {}
    for s, lab in enumerate(labels):
        for o, c in enumerate(lab):
            if o == 0:
                first_row[s] = len(rows)
            rows.append(c)
            preds.append([len(rows) - 2] if o > 0 else None)
            last_row[s] = len(rows) - 1
    has_pred, has_succ = set(), set()
    pl = {s: [] for s in range(len(labels))}
    for a, b in edges:
        pl[b].append(a); has_pred.add(b); has_succ.add(a)
    for s in range(len(labels)):
        preds[first_row[s]] = [last_row[a] for a in pl[s]] if s in has_pred else [-1]
    n = len(q)
    oe1, oe2 = P.o1 + P.e1, P.o2 + P.e2
    srcH = [0] + [-gap(j, P) for j in range(1, n + 1)]
    H, E1, E2 = {}, {}, {}
    H[-1] = srcH
    E1[-1] = [x - oe1 for x in srcH]
    E2[-1] = [x - oe2 for x in srcH]
    for i, c in enumerate(rows):
        h = [NEG] * (n + 1); e1 = [NEG] * (n + 1); e2 = [NEG] * (n + 1)
        f1 = f2 = NEG
        for j in range(n + 1):
            m = max((H[p][j - 1] for p in preds[i]), default=NEG) + (P.match if q[j - 1] == c else -P.mismatch) if j else NEG
            x1 = max(E1[p][j] for p in preds[i]); x2 = max(E2[p][j] for p in preds[i])
            if j:
                f1 = max(h[j - 1] - oe1, f1 - P.e1); f2 = max(h[j - 1] - oe2, f2 - P.e2)
            h[j] = max(m, x1, x2, f1, f2)
            e1[j] = max(x1 - P.e1, h[j] - oe1); e2[j] = max(x2 - P.e2, h[j] - oe2)
        H[i], E1[i], E2[i] = h, e1, e2
    return max(H[last_row[s]][n] for s in range(len(labels)) if s not in has_succ)


def random_case(rng, n_sites, err):
    """A bubble chain (node 0 has no predecessor) and a read walked through it, with errors."""
    labels, edges, tails = [], [], []
    for _ in range(n_sites):
        labels.append("".join(rng.choice("ACGT") for _ in range(rng.randint(1, 12))))
        cur = len(labels) - 1
        edges += [(t, cur) for t in tails]
        kind = rng.random()
        if kind < 0.5:
            a = rng.choice("ACGT"); b = rng.choice([c for c in "ACGT" if c != a])
            labels += [a, b]
            edges += [(cur, cur + 1), (cur, cur + 2)]
            tails = [cur + 1, cur + 2]
        elif kind < 0.8:
            labels.append("".join(rng.choice("ACGT") for _ in range(rng.randint(1, 6))))
            edges.append((cur, cur + 1))
            tails = [cur, cur + 1]
        else:
            tails = [cur]
    labels.append("".join(rng.choice("ACGT") for _ in range(rng.randint(2, 10))))
    edges += [(t, len(labels) - 1) for t in tails]
    edges.sort()
    succ = {}
    for a, b in edges:
        succ.setdefault(a, []).append(b)
    s, walk = 0, ""
    while True:
        walk += labels[s]
        if s not in succ:
            break
        s = rng.choice(succ[s])
    read = ""
    for c in walk:
        u = rng.random()
        if u < err / 3:
            continue
        if u < 2 * err / 3:
            read += rng.choice("ACGT")
        if u < err:
            read += rng.choice([x for x in "ACGT" if x != c])
        else:
            read += c
    return labels, edges, read or "A"


def test_exact_match_linear():
    P = orc.default_params()
    r = orc.poa_align(["ACGT", "TTGA", "C"], [(0, 1), (1, 2)], "ACGTTTGAC", P)
    assert r.aln_status[0] == 1 and r.aln_score[0] == 18
    assert runs_of(r) == [(M, 9)]
    assert r.aln_path.tolist() == [0, 1, 2] and r.aln_path_len[0] == 9
    assert (r.aln_start_off[0], r.aln_end_off[0], r.aln_n_aligned[0]) == (0, 0, 9)


def test_snp_bubble_takes_the_matching_allele():
    P = orc.default_params()
    labels, edges = ["ACGTAC", "A", "G", "TTGACC"], [(0, 1), (0, 2), (1, 3), (2, 3)]
    assert orc.poa_align(labels, edges, "ACGTACGTTGACC", P).aln_path.tolist() == [0, 2, 3]
    assert orc.poa_align(labels, edges, "ACGTACATTGACC", P).aln_path.tolist() == [0, 1, 3]
    r = orc.poa_align(labels, edges, "ACGTACCTTGACC", P)      # mismatch against both: first predecessor wins
    assert r.aln_path.tolist() == [0, 1, 3] and runs_of(r) == [(M, 6), (X, 1), (M, 6)]


def test_insertion_deletion_ops():
    P = orc.default_params()
    r = orc.poa_align(["ACGTACGTAC", "GGTTGGTTGG"], [(0, 1)], "ACGTACGTACAAAGGTTGGTTGG", P)
    assert runs_of(r) == [(M, 10), (I, 3), (M, 10)] and r.aln_score[0] == 40 - 10
    r = orc.poa_align(["ACGTACGTAC", "TTT", "GGTTGGTTGG"], [(0, 1), (1, 2)], "ACGTACGTACGGTTGGTTGG", P)
    assert runs_of(r) == [(M, 10), (D, 3), (M, 10)] and r.aln_path_len[0] == 23


def test_leading_insertion_and_no_leading_deletion():
    P = orc.default_params()
    r = orc.poa_align(["ACGTACGTAC"], [], "TTACGTACGTAC", P)
    assert runs_of(r) == [(I, 2), (M, 10)]
    r = orc.poa_align(["GGGGGACGTACGTAC"], [], "ACGTACGTAC", P)     # leading graph bases are not reported
    assert runs_of(r) == [(M, 10)] and r.aln_start_off[0] == 5 and r.aln_end_off[0] == 14
    assert r.aln_score[0] == 20 - min(4 + 2 * 5, 24 + 5)


@pytest.mark.parametrize("seed", range(40))
def test_score_matches_unbanded_dp_and_traceback_rescoring(seed):
    rng = random.Random(seed)
    labels, edges, read = random_case(rng, rng.randint(1, 8), rng.choice([0.0, 0.05, 0.15, 0.3]))
    wide = orc.default_params(wb=len(read) + sum(map(len, labels)))     # band = the whole matrix
    r = orc.poa_align(labels, edges, read, wide)
    assert r.aln_status[0] == 1
    assert r.aln_score[0] == brute_force(labels, edges, read, wide)
    runs = runs_of(r)
    assert sum(n for op, n in runs if op != D) == len(read)
    if r.aln_start_off[0] == 0 and int(r.aln_path[0]) == 0:
        assert rescore(runs, wide) == r.aln_score[0]
    # the default adaptive band may only lose score, never invent it
    rb = orc.poa_align(labels, edges, read, orc.default_params())
    assert rb.aln_status[0] in (1, 2)
    if rb.aln_status[0] == 1:
        assert rb.aln_score[0] <= r.aln_score[0]
        assert rb.aln_cells[0] <= r.aln_cells[0]


# ---- abPOA-compatibility switches: constants a later comparison with real abPOA flips
def _rand_case(rng, n_nodes=12, qlen=60):
    labels = ["".join(rng.choice(list("ACGT"), size=int(rng.integers(1, 9)))) for _ in range(n_nodes)]
    edges = [(i, i + 1) for i in range(n_nodes - 1)] + [(i, i + 2) for i in range(0, n_nodes - 2, 3)]
    edges.sort()
    q = "".join(rng.choice(list("ACGT"), size=qlen))
    return labels, edges, q


def test_band_granularity_widens_the_band_and_never_lowers_the_score():
    """ORC_FLAG_BAND_GRAN8 / 16: every row's band is widened to vector boundaries, so the cells of granularity 1 are a subset
    (cell counts grow, the score cannot drop), and with a band that already spans the read all three agree."""
    rng = np.random.default_rng(5)
    grew = 0
    for _ in range(30):
        labels, edges, q = _rand_case(rng)
        r1 = orc.poa_align(labels, edges, q)
        r8 = orc.poa_align(labels, edges, q, orc.default_params(flags=orc.FLAG_BAND_GRAN8))
        r16 = orc.poa_align(labels, edges, q, orc.default_params(flags=orc.FLAG_BAND_GRAN16))
        if r1.aln_status[0] != 1:
            continue
        assert r8.aln_status[0] == 1 and r16.aln_status[0] == 1
        assert r1.aln_cells[0] <= r8.aln_cells[0] <= r16.aln_cells[0]
        assert r1.aln_score[0] <= r8.aln_score[0] <= r16.aln_score[0]
        grew += int(r16.aln_cells[0] > r1.aln_cells[0])
        wide = orc.default_params(wb=1000)
        w8 = orc.default_params(wb=1000, flags=orc.FLAG_BAND_GRAN8)
        a, b = orc.poa_align(labels, edges, q, wide), orc.poa_align(labels, edges, q, w8)
        assert a.aln_score[0] == b.aln_score[0] and np.array_equal(a.aln_runs, b.aln_runs)
    assert grew > 10


def test_int16_sentinel_changes_nothing_while_scores_are_far_from_it():
    rng = np.random.default_rng(6)
    for _ in range(20):
        labels, edges, q = _rand_case(rng)
        a = orc.poa_align(labels, edges, q)
        b = orc.poa_align(labels, edges, q, orc.default_params(flags=orc.FLAG_INT16_SCORES))
        for f in ("aln_status", "aln_score", "aln_runs", "aln_path", "aln_cells"):
            assert np.array_equal(getattr(a, f), getattr(b, f)), f
