// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See oracle.h for the usage rule and parity status.
//
// CPU restatement of the reference's `map --also-align` path.  Each function cites the reference
// file:line it follows (paths relative to the reference tree).  Results-identical replacements of
// the reference's O(N) scans (hash map for `kmer_pos_ref.contains`, binary search for rank/select,
// no sequence clones) are marked "sane:".
//
// Build: g++ -O2 -std=c++17 -ffp-contract=off -pthread -shared -fPIC  (Rust never contracts a*b+c
// into an FMA, so neither may this file).
#include "oracle.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <deque>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>
#include <atomic>
#include <thread>

namespace {

typedef uint64_t u64;
typedef uint32_t u32;
const u64 REV_BIT = 1ull << 63;
const u64 DELIM = ~0ull;

// ---------------------------------------------------------------------------------------------
// Index (src/index.rs:36-90)
// ---------------------------------------------------------------------------------------------
struct Index {
    u32 k;
    u64 L, n_nodes, n_edges;
    std::string fwd, rev;
    std::vector<u64> seq_idx, edge_idx, edges_in;  // n_nodes+1 each (NodeRef, src/utils.rs:14-22)
    std::vector<u64> edges;
    std::unordered_map<std::string, u64> key2off;  // sane: stands for kmer_pos_ref.contains + bhpf.get
    std::vector<u64> tab_start, tab_end;

    // get_bv_rank (src/index.rs:427-439): set bits of seq_bv in [0, pos]; bits are set at every
    // node start and at L (src/utils.rs:130,135).
    u64 rank(u64 pos) const {
        u64 r = std::upper_bound(seq_idx.begin(), seq_idx.begin() + n_nodes, pos) - seq_idx.begin();
        if (pos >= L) r += 1;
        return r;
    }
    // get_bv_inverse_rank (src/index.rs:443-458): set bits in [L-pos, L]
    u64 inverse_rank(u64 pos) const {
        u64 lo = L - pos;
        u64 below = std::lower_bound(seq_idx.begin(), seq_idx.begin() + n_nodes, lo) - seq_idx.begin();
        return (n_nodes - below) + 1;
    }
    // get_bv_select (src/index.rs:461-480): position of the element_no-th set bit, 0 if absent
    u64 select(u64 element_no) const {
        if (element_no >= 1 && element_no <= n_nodes) return seq_idx[element_no - 1];
        if (element_no == n_nodes + 1) return L;
        return 0;
    }
    // node_id_from_seqpos / handle_from_seqpos (src/index.rs:388-423)
    u64 handle_from_seqpos(bool reverse, u64 pos) const {
        if (!reverse) return rank(pos) * 2;
        return (n_nodes - inverse_rank(pos) + 1) * 2 + 1;
    }
    u64 node_len(u64 id) const { return seq_idx[id] - seq_idx[id - 1]; }
    // incoming_edges_from_handle / outgoing_edges_from_handle (src/index.rs:559-606)
    std::vector<u64> incoming(u64 h) const {
        u64 id = h >> 1;
        if (!(h & 1)) {
            u64 b = edge_idx[id - 1];
            return std::vector<u64>(edges.begin() + b, edges.begin() + b + edges_in[id - 1]);
        }
        std::vector<u64> o = outgoing(h ^ 1);
        for (auto& x : o) x ^= 1;
        std::reverse(o.begin(), o.end());
        return o;
    }
    std::vector<u64> outgoing(u64 h) const {
        u64 id = h >> 1;
        if (!(h & 1)) {
            u64 b = edge_idx[id - 1] + edges_in[id - 1], e = edge_idx[id];
            return std::vector<u64>(edges.begin() + b, edges.begin() + e);
        }
        std::vector<u64> o = incoming(h ^ 1);
        for (auto& x : o) x ^= 1;
        std::reverse(o.begin(), o.end());
        return o;
    }
    // seq_from_handle (src/index.rs:503-533)
    std::string seq_from_handle(u64 h) const {
        u64 id = h >> 1;
        if (!(h & 1)) return fwd.substr(seq_idx[id - 1], seq_idx[id] - seq_idx[id - 1]);
        return rev.substr(L - seq_idx[id], seq_idx[id] - seq_idx[id - 1]);
    }
};

// reverse_complement (src/dna.rs:5-33)
std::string revcomp(const std::string& s) {
    std::string o(s.size(), 'N');
    for (size_t i = 0; i < s.size(); i++) {
        char c = s[s.size() - 1 - i], r = 'N';
        switch (c) {
            case 'a': r = 't'; break; case 'c': r = 'g'; break; case 't': r = 'a'; break;
            case 'g': r = 'c'; break; case 'u': r = 'a'; break; case 'A': r = 'T'; break;
            case 'C': r = 'G'; break; case 'T': r = 'A'; break; case 'G': r = 'C'; break;
            case 'U': r = 'A'; break; default: r = 'N';
        }
        o[i] = r;
    }
    return o;
}

// ---------------------------------------------------------------------------------------------
// Seeding (src/chain.rs:134-173, src/index.rs:309-382, src/io.rs:41-56)
// ---------------------------------------------------------------------------------------------
struct Anchor {
    u64 id, qb, qe;
    u64 tb, te;          // positions
    bool tb_rev, te_rev;  // orientations
    double f;            // max_chain_score
    int64_t pred;        // best_predecessor_id, -1 = None
};

// find_positions_for_query_kmer: entries from the start offset up to the delimiter
void positions_for(const Index& ix, const char* kmer, u64 len, std::vector<u64>& s, std::vector<u64>& e) {
    s.clear(); e.clear();
    if (len != ix.k) return;                                   // src/index.rs:310-312
    auto it = ix.key2off.find(std::string(kmer, len));         // :319-320
    if (it == ix.key2off.end()) return;
    u64 p = it->second;
    while (!(ix.tab_start[p] == DELIM && ix.tab_end[p] == DELIM)) {   // :328-348
        s.push_back(ix.tab_start[p]);
        e.push_back(ix.tab_end[p]);
        p++;
    }
}

void anchors_for_query(const Index& ix, const char* read, u64 len, bool only_forward, std::vector<Anchor>& out) {
    out.clear();
    u64 k = ix.k;
    if (k > len) return;                                        // src/io.rs:47
    u64 id = 0;
    std::vector<u64> ps, pe;
    for (u64 i = 0; i + k <= len; i++) {                        // src/chain.rs:148
        positions_for(ix, read + i, k, ps, pe);
        for (size_t t = 0; t < ps.size(); t++) {
            bool srev = ps[t] & REV_BIT, erev = pe[t] & REV_BIT;
            if ((only_forward && !srev && !erev) || !only_forward) {      // :154-156
                Anchor a{id, i, i + k, ps[t] & ~REV_BIT, pe[t] & ~REV_BIT, srev, erev, (double)k, -1};
                out.push_back(a);
                id++;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Chaining (src/chain.rs:274-655)
// ---------------------------------------------------------------------------------------------
double score_anchor(const Anchor& a, const Anchor& b, u64 seed_length, u64 max_gap) {
    double score = a.f;
    if (a.qe >= b.qe || (a.te_rev == b.te_rev && a.te >= b.te) ||
        !(a.te_rev == b.te_rev && a.tb_rev == b.tb_rev && a.te_rev == b.tb_rev && a.tb_rev == b.te_rev)) {
        score = -std::numeric_limits<double>::max();                      // :277-311
    } else {
        u64 query_length = std::min(b.qb - a.qb, b.qe - a.qe);             // :313
        u64 query_overlap = a.qe > b.qe ? a.qe - b.qe : 0;                 // :315-318 (always 0 here)
        // SeqPos comparison is (orient, position) (src/kmer.rs:27-31)
        auto gt = [](bool r1, u64 p1, bool r2, u64 p2) { return r1 != r2 ? (r1 > r2) : (p1 > p2); };
        u64 tbd = gt(b.tb_rev, b.tb, a.tb_rev, a.tb) ? b.tb - a.tb : a.tb - b.tb;   // :324-327
        u64 ted = gt(b.te_rev, b.te, a.te_rev, a.te) ? b.te - a.te : a.te - b.te;   // :328-331
        u64 target_length = std::min(tbd, ted);
        u64 gap_length = query_length > target_length ? query_length - target_length
                                                      : target_length - query_length;   // :338-341
        if (gap_length > max_gap) {
            score = -std::numeric_limits<double>::max();
        } else {
            double gap_cost = gap_length == 0 ? 0.0
                                              : 0.01 * (double)seed_length * (double)gap_length +
                                                    0.5 * std::log2((double)gap_length);   // :348-354
            u64 match_length = std::min(std::min(query_length, target_length), seed_length);
            score = std::round((a.f + (double)match_length - gap_cost) * 1000.0) / 1000.0 +
                    (double)query_overlap;                                              // :360-363
        }
    }
    return score;
}

struct Chain {
    std::vector<Anchor> anchors;  // query order
    bool placeholder = false;
};

struct ChainDebug {
    std::vector<u32> sorted;   // anchor id by sorted position
    std::vector<double> f;     // by sorted position
    std::vector<int32_t> pred; // pred anchor id by sorted position, before backtracking consumes it
};

void chain_anchors(std::vector<Anchor>& anchors, u64 seed_length, u64 bandwidth, u64 max_gap, u64 min_n,
                   std::vector<Chain>& chains, ChainDebug* dbg) {
    chains.clear();
    // :386-389 stable sort, orientation descending then target_end.position ascending
    std::stable_sort(anchors.begin(), anchors.end(), [](const Anchor& a, const Anchor& b) {
        if (a.te_rev != b.te_rev) return a.te_rev > b.te_rev;
        return a.te < b.te;
    });
    double curr_max = 0.0;
    const size_t n = anchors.size();
    for (size_t i = 1; i < n; i++) {                                       // :403-450
        size_t min_j = bandwidth > i ? 0 : i - bandwidth;
        for (size_t j = i; j-- > min_j;) {
            double s = score_anchor(anchors[j], anchors[i], seed_length, max_gap);
            if (s > anchors[i].f) {
                anchors[i].f = s;
                anchors[i].pred = (int64_t)anchors[j].id;
            }
            if (s > curr_max) curr_max = s;
        }
    }
    if (dbg) {
        dbg->sorted.resize(n); dbg->f.resize(n); dbg->pred.resize(n);
        for (size_t i = 0; i < n; i++) {
            dbg->sorted[i] = (u32)anchors[i].id;
            dbg->f[i] = anchors[i].f;
            dbg->pred[i] = (int32_t)anchors[i].pred;
        }
    }
    if (n > 0) {                                                           // :455-558
        std::vector<size_t> pos_of_id(n);   // sane: replaces anchors.iter().position(id) (:493)
        for (size_t i = 0; i < n; i++) pos_of_id[anchors[i].id] = i;
        for (size_t i = n; i-- > 0;) {
            size_t cur = i;
            if (anchors[cur].pred >= 0 && anchors[cur].f == curr_max) {     // :469
                Chain c;
                while (anchors[cur].pred >= 0) {
                    int64_t p = anchors[cur].pred;
                    anchors[cur].pred = -1;                                 // consumes the predecessor link
                    c.anchors.push_back(anchors[cur]);
                    cur = pos_of_id[(size_t)p];
                }
                c.anchors.push_back(anchors[cur]);
                if (c.anchors.size() >= min_n) {                            // :545-548
                    std::reverse(c.anchors.begin(), c.anchors.end());
                    chains.push_back(std::move(c));
                }
            }
        }
    }
    // :563 sort by score desc — every Chain.score is 0.0, a stable no-op
    if (chains.empty()) {                                                   // :644-649
        Chain c;
        c.placeholder = true;
        chains.push_back(c);
    }
}

// ---------------------------------------------------------------------------------------------
// Chain -> subgraph (src/align.rs:267-402, 523-665, 670-724)
// ---------------------------------------------------------------------------------------------
enum { ST_PLACEHOLDER = 0, ST_OK = 1, ST_EMPTY = 2, ST_TRACE_ERR = 3, ST_BAND_ERR = 4, ST_UNSUPPORTED = 5 };

bool find_range_chain(const Index& ix, const Chain& c, std::vector<u64>& handles) {
    handles.clear();
    u64 mn = ~0ull, mx = 0;
    for (const Anchor& a : c.anchors) {
        u64 h1 = ix.handle_from_seqpos(a.tb_rev, a.tb);                     // :286-290
        u64 h2 = ix.handle_from_seqpos(a.te_rev, a.te - 1);                 // :291-298 (inclusive end, chain.rs:65-70)
        mn = std::min(mn, std::min(h1, h2));
        mx = std::max(mx, std::max(h1, h2));
    }
    if ((mn & 1) || (mx & 1)) return false;   // Reverse / Both ranges: unreachable with only_forward=true
    // :358-364.  u64::from(Handle) is taken as the node id.
    for (u64 x = mn >> 1; x <= (mx >> 1); x++) handles.push_back(x * 2);
    return true;
}

// extend_range_chain_2 (src/align.rs:523-665).  `literal` walks the frontier exactly as written
// (no visited set: exponential on bubble chains); otherwise a node is re-expanded only when it
// is reached with a larger remaining length, which yields the same handle SET (the remaining
// length only ever widens what gets pushed), and the set is all the function returns
// (sort + dedup at :658-659).
void extend_range(const Index& ix, const Chain& c, u64 qlen, std::vector<u64>& handles, bool literal) {
    const u64 first = handles.front(), last = handles.back();
    const Anchor& a0 = c.anchors.front();
    const Anchor& a1 = c.anchors.back();
    std::vector<u64> ext;
    auto walk = [&](u64 start, u64 budget, bool left) {
        std::vector<std::pair<u64, u64>> frontier, next;
        std::unordered_map<u64, u64> best;
        for (u64 h : (left ? ix.incoming(start) : ix.outgoing(start))) frontier.push_back({budget, h});
        while (!frontier.empty()) {
            next.clear();
            for (auto& pr : frontier) {
                u64 rem = pr.first, h = pr.second;
                if (!literal) {
                    auto it = best.find(h);
                    if (it != best.end() && it->second >= rem) continue;
                    best[h] = rem;
                }
                ext.push_back(h);
                u64 len = ix.node_len(h >> 1);
                if (len < rem)
                    for (u64 x : (left ? ix.incoming(h) : ix.outgoing(h))) next.push_back({rem - len, x});
            }
            frontier.swap(next);
        }
    };
    u64 prefix = a0.qb;                                                      // :536
    u64 slack = a0.tb - ix.select(first >> 1);                               // :541-542
    prefix = slack < prefix ? prefix - slack : 0;                            // :543-547
    if (prefix > 0) walk(first, prefix, true);                               // :551-591
    u64 suffix = qlen - a1.qe;                                               // :593-594
    u64 eslack = ix.select((last >> 1) + 1) - 1 - (a1.te - 1);               // :599-601
    suffix = eslack > suffix ? 0 : suffix - eslack;                          // :608-612
    if (suffix > 0) walk(last, suffix, false);                               // :616-656
    handles.insert(handles.end(), ext.begin(), ext.end());
    std::sort(handles.begin(), handles.end());                               // :658-659
    handles.erase(std::unique(handles.begin(), handles.end()), handles.end());
}

// find_nodes_edges_for_abpoa (src/align.rs:670-724), Forward range
void nodes_edges(const Index& ix, const std::vector<u64>& handles, std::vector<std::string>& labels,
                 std::vector<std::pair<u32, u32>>& edges) {
    labels.clear(); edges.clear();
    for (u64 h : handles) labels.push_back(ix.seq_from_handle(h));
    for (size_t s = 0; s < handles.size(); s++) {
        for (u64 t : ix.outgoing(handles[s])) {
            auto it = std::lower_bound(handles.begin(), handles.end(), t);   // sane: replaces contains/position
            if (it == handles.end() || *it != t) continue;
            size_t e = it - handles.begin();
            if (s < e) edges.push_back({(u32)s, (u32)e});                    // :717-718 keep s<e
        }
    }
}

// ---------------------------------------------------------------------------------------------
// POA (stands for ab_poa::AbpoaAligner::create_align_safe, src/align.rs:190-202).  PARITY UNPINNED.
// Global alignment, convex gap min(o1+g*e1, o2+g*e2), adaptive band w = wb + (int)(wf*qlen);
// abPOA's recurrences and value-driven backtrack.  Every tie-break below is part of the definition
// the CUDA path must reproduce.
// ---------------------------------------------------------------------------------------------
const int32_t NEG_INF_32 = -(1 << 29);
// ORC_BAND_DUMP=<file>: one line per DP row (row, beg, end, #pred rows, pred is the previous row, qlen) — band statistics for
// kernel design; single-threaded runs only
static FILE* g_band_dump = getenv("ORC_BAND_DUMP") ? fopen(getenv("ORC_BAND_DUMP"), "w") : nullptr;

struct AlnResult {
    uint8_t status = ST_PLACEHOLDER;
    int32_t score = 0;
    u64 cells = 0;
    std::vector<u32> runs;       // op<<28 | len ; op: 0 '=', 1 'X', 2 'D', 3 'I'
    std::vector<u32> path_nodes; // subgraph node indices after dedup
    u32 path_len = 0, start_off = 0, end_off = 0, n_aligned = 0;
};

struct Row {
    int32_t beg, end;
    std::vector<int32_t> H, E1, E2, F1, F2;   // E1/E2 are the values handed to successor rows (E')
};

void poa_align(const std::vector<std::string>& labels, const std::vector<std::pair<u32, u32>>& edges,
               const char* q, int32_t qlen, const orc_params_t& P, AlnResult& R) {
    const int32_t S = (int32_t)labels.size();
    const int32_t oe1 = P.o1 + P.e1, oe2 = P.o2 + P.e2, e1 = P.e1, e2 = P.e2;
    const int32_t w = P.wb + (int32_t)(P.wf * (float)qlen);
    // compatibility switches (oracle.h): score width of the -inf sentinel, band granularity
    const int32_t NEG_INF = (P.flags & ORC_FLAG_INT16_SCORES) ? INT16_MIN + std::max(P.mismatch, std::max(oe1, oe2)) : NEG_INF_32;
    const int32_t gmask = (P.flags & ORC_FLAG_BAND_GRAN16) ? 15 : (P.flags & ORC_FLAG_BAND_GRAN8) ? 7 : 0;
    std::vector<int32_t> row_start(S + 1, 0);
    for (int32_t s = 0; s < S; s++) row_start[s + 1] = row_start[s] + (int32_t)labels[s].size();
    const int32_t nrows = row_start[S];
    // node-level predecessor / successor lists in edge-list order
    std::vector<std::vector<int32_t>> preds(S), succs(S);
    for (auto& e : edges) { preds[e.second].push_back((int32_t)e.first); succs[e.first].push_back((int32_t)e.second); }
    // max_remain along the first out-edge (abPOA: heaviest out-edge, all weights equal -> first)
    std::vector<int32_t> remain_first(S), remain_last(S);
    for (int32_t s = S - 1; s >= 0; s--) {
        remain_last[s] = succs[s].empty() ? 0 : remain_first[succs[s][0]] + 1;
        remain_first[s] = remain_last[s] + (int32_t)labels[s].size() - 1;
    }
    const int32_t remain_src = remain_first[0] + 1;
    // rows: DP index 0 = source, 1..nrows = bases
    std::vector<Row> rows(nrows + 1);
    std::vector<int32_t> mpl(nrows + 1, INT_MAX), mpr(nrows + 1, 0);
    {   // source row
        Row& r = rows[0];
        r.beg = 0;
        r.end = std::min(qlen, std::max(0, qlen - remain_src) + w);
        r.end = std::min(qlen, r.end | gmask);                       // band granularity (g = 1: unchanged)
        int32_t W = r.end + 1;
        r.H.assign(W, NEG_INF); r.E1.assign(W, NEG_INF); r.E2.assign(W, NEG_INF);
        r.H[0] = 0; r.E1[0] = -oe1; r.E2[0] = -oe2;
        for (int32_t j = 1; j <= r.end; j++) r.H[j] = std::max(-(P.o1 + P.e1 * j), -(P.o2 + P.e2 * j));
        for (int32_t s = 0; s < S; s++)
            if (preds[s].empty()) { mpl[row_start[s] + 1] = 1; mpr[row_start[s] + 1] = 1; }
    }   // the source row is not counted in `cells` (cells are summed over subgraph bases)
    auto pred_rows = [&](int32_t s, int32_t o, std::vector<int32_t>& out) {
        out.clear();
        if (o > 0) { out.push_back(row_start[s] + o); return; }            // previous base of the node
        if (preds[s].empty()) { out.push_back(0); return; }                // virtual source
        for (int32_t p : preds[s]) out.push_back(row_start[p + 1]);         // last base of each pred node
    };
    std::vector<int32_t> pr;
    for (int32_t s = 0; s < S; s++) {
        const int32_t len = (int32_t)labels[s].size();
        for (int32_t o = 0; o < len; o++) {
            const int32_t i = row_start[s] + o + 1;
            const char base = labels[s][o];
            pred_rows(s, o, pr);
            const int32_t remain = remain_last[s] + (len - 1 - o);
            const int32_t r = qlen - remain;
            int32_t beg = std::max(0, std::min(mpl[i], r) - w);
            int32_t end = std::min(qlen, std::max(mpr[i], r) + w);
            int32_t min_pb = INT_MAX, max_pe = -1;
            for (int32_t p : pr) { min_pb = std::min(min_pb, rows[p].beg); max_pe = std::max(max_pe, rows[p].end); }
            beg = std::max(beg, min_pb);
            end = std::min(end, max_pe + 1);
            // band granularity: widened to vector boundaries, clipped again on the right (every begin is a multiple of g,
            // so the widened begin stays inside the predecessors' bands)
            beg &= ~gmask;
            end = std::min(std::min(qlen, end | gmask), max_pe + 1);
            if (beg > end) { R.status = ST_BAND_ERR; return; }
            Row& cur = rows[i];
            cur.beg = beg; cur.end = end;
            const int32_t W = end - beg + 1;
            cur.H.assign(W, NEG_INF); cur.E1.assign(W, NEG_INF); cur.E2.assign(W, NEG_INF);
            cur.F1.assign(W, NEG_INF); cur.F2.assign(W, NEG_INF);
            std::vector<int32_t> hhat(W);
            for (int32_t j = beg; j <= end; j++) {
                int32_t hd = NEG_INF, x1 = NEG_INF, x2 = NEG_INF;
                for (int32_t p : pr) {
                    const Row& pw = rows[p];
                    if (j - 1 >= pw.beg && j - 1 <= pw.end) hd = std::max(hd, pw.H[j - 1 - pw.beg]);
                    if (j >= pw.beg && j <= pw.end) {
                        x1 = std::max(x1, pw.E1[j - pw.beg]);
                        x2 = std::max(x2, pw.E2[j - pw.beg]);
                    }
                }
                int32_t m = j >= 1 ? hd + (q[j - 1] == base ? P.match : -P.mismatch) : NEG_INF;
                cur.E1[j - beg] = x1; cur.E2[j - beg] = x2;   // E(i,j) for now
                hhat[j - beg] = std::max(m, std::max(x1, x2));
            }
            // F(beg) = -inf; F(beg+1) = Hhat(beg)-oe; F(j) = max(Hhat(j-1)-oe, F(j-1)-e)
            for (int32_t t = 1; t < W; t++) {
                cur.F1[t] = t == 1 ? hhat[0] - oe1 : std::max(hhat[t - 1] - oe1, cur.F1[t - 1] - e1);
                cur.F2[t] = t == 1 ? hhat[0] - oe2 : std::max(hhat[t - 1] - oe2, cur.F2[t - 1] - e2);
            }
            int32_t mx = INT_MIN, lpos = beg, rpos = beg;
            for (int32_t t = 0; t < W; t++) {
                int32_t h = std::max(hhat[t], std::max(cur.F1[t], cur.F2[t]));
                cur.H[t] = h;
                cur.E1[t] = std::max(cur.E1[t] - e1, h - oe1);   // E' handed to successors
                cur.E2[t] = std::max(cur.E2[t] - e2, h - oe2);
                if (h > mx) { mx = h; lpos = beg + t; }
                if (h >= mx) rpos = beg + t;
            }
            R.cells += W;
            if (g_band_dump) fprintf(g_band_dump, "%d %d %d %d %d %d\n", i, beg, end, (int)pr.size(), pr[0] == i - 1 ? 1 : 0, qlen);
            // adaptive band hints for the out rows (abPOA simd_abpoa_ada_max_i)
            auto hint = [&](int32_t t) {
                mpr[t] = std::max(mpr[t], rpos + 1);
                mpl[t] = std::min(mpl[t], lpos + 1);
            };
            if (o + 1 < len) hint(i + 1);
            else for (int32_t t : succs[s]) hint(row_start[t] + 1);
        }
    }
    // global end: sink predecessors (nodes without successors, ascending) whose band reaches qlen
    int32_t best = INT_MIN, bi = -1, bs = -1;
    for (int32_t s = 0; s < S; s++) {
        if (!succs[s].empty()) continue;
        int32_t i = row_start[s + 1];
        if (rows[i].end < qlen) continue;
        int32_t h = rows[i].H[qlen - rows[i].beg];
        if (h > best) { best = h; bi = i; bs = s; }
    }
    if (bi < 0) { R.status = ST_EMPTY; return; }
    R.score = best;
    // ---- backtrack
    enum { M_OP = 1, E1_OP = 2, E2_OP = 4, F1_OP = 8, F2_OP = 16, ALL_OP = 31 };
    struct Op { uint8_t op; int32_t row; };
    std::vector<Op> ops;
    int32_t i = bi, j = qlen, s = bs, o = (int32_t)labels[bs].size() - 1, cur = ALL_OP;
    auto inband = [&](int32_t p, int32_t jj) { return jj >= rows[p].beg && jj <= rows[p].end; };
    while (i != 0 && j > 0) {
        pred_rows(s, o, pr);
        const Row& ri = rows[i];
        const int32_t sc = q[j - 1] == labels[s][o] ? P.match : -P.mismatch;
        const int32_t hij = ri.H[j - ri.beg];
        bool hit = false;
        int32_t chosen = -1;     // predecessor list index
        for (int attempt = 0; attempt < 2 && !hit; attempt++) {
            if (cur & M_OP) {
                for (size_t kx = 0; kx < pr.size(); kx++) {
                    int32_t p = pr[kx];
                    if (!inband(p, j - 1)) continue;
                    if (rows[p].H[j - 1 - rows[p].beg] + sc == hij) {
                        ops.push_back({(uint8_t)(q[j - 1] == labels[s][o] ? 0 : 1), i});
                        chosen = (int32_t)kx; j--; cur = ALL_OP; hit = true;
                        break;
                    }
                }
            }
            if (!hit && (cur & (E1_OP | E2_OP))) {
                for (size_t kx = 0; kx < pr.size() && !hit; kx++) {
                    int32_t p = pr[kx];
                    if (!inband(p, j)) continue;
                    const Row& rp = rows[p];
                    int32_t pe1 = rp.E1[j - rp.beg], pe2 = rp.E2[j - rp.beg], ph = rp.H[j - rp.beg];
                    if (cur & E1_OP) {
                        bool c = (cur & M_OP) ? (hij == pe1) : (ri.E1[j - ri.beg] == pe1 - e1);
                        if (c) { cur = (ph - oe1 == pe1) ? (M_OP | F1_OP | F2_OP) : E1_OP; hit = true; }
                    }
                    if (!hit && (cur & E2_OP)) {
                        bool c = (cur & M_OP) ? (hij == pe2) : (ri.E2[j - ri.beg] == pe2 - e2);
                        if (c) { cur = (ph - oe2 == pe2) ? (M_OP | F1_OP | F2_OP) : E2_OP; hit = true; }
                    }
                    if (hit) { ops.push_back({2, i}); chosen = (int32_t)kx; }
                }
            }
            if (!hit && (cur & (F1_OP | F2_OP)) && j > ri.beg) {
                int32_t hl = ri.H[j - 1 - ri.beg];
                if (cur & F1_OP) {
                    int32_t f = ri.F1[j - ri.beg];
                    if (!(cur & M_OP) || hij == f) { cur = (hl - oe1 == f) ? (M_OP | E1_OP | E2_OP) : F1_OP; hit = true; }
                }
                if (!hit && (cur & F2_OP)) {
                    int32_t f = ri.F2[j - ri.beg];
                    if (!(cur & M_OP) || hij == f) { cur = (hl - oe2 == f) ? (M_OP | E1_OP | E2_OP) : F2_OP; hit = true; }
                }
                if (hit) { ops.push_back({3, -1}); j--; }
            }
            if (!hit) {
                if (cur == ALL_OP) break;
                cur = ALL_OP;    // definition: an exhausted restricted state falls back to ALL once
            }
        }
        if (!hit) { R.status = ST_TRACE_ERR; return; }
        if (chosen >= 0) {       // moved to a predecessor row
            if (o > 0) { o--; i--; }
            else if (preds[s].empty()) { i = 0; }
            else { s = preds[s][chosen]; o = (int32_t)labels[s].size() - 1; i = row_start[s + 1]; }
        }
    }
    if (i == 0) for (; j > 0; j--) ops.push_back({3, -1});   // leading insertions; no leading deletions
    std::reverse(ops.begin(), ops.end());
    // ---- derived outputs (rs-abPOA's AbpoaAlignmentResult fields, as defined in DESIGN.md)
    int32_t last_node = -1, first_row = -1, last_row = -1;
    for (const Op& op : ops) {
        if (!R.runs.empty() && (R.runs.back() >> 28) == op.op) R.runs.back() += 1;
        else R.runs.push_back(((u32)op.op << 28) | 1u);
        if (op.op <= 1) R.n_aligned++;
        if (op.op <= 2) {
            R.path_len++;
            int32_t row0 = op.row - 1;
            int32_t node = (int32_t)(std::upper_bound(row_start.begin(), row_start.end(), row0) - row_start.begin()) - 1;
            if (node != last_node) { R.path_nodes.push_back((u32)node); last_node = node; }
            if (first_row < 0) first_row = row0;
            last_row = row0;
        }
    }
    if (first_row >= 0) {
        R.start_off = (u32)(first_row - row_start[R.path_nodes.front()]);
        R.end_off = (u32)(last_row - row_start[R.path_nodes.back()]);
    }
    R.status = ST_OK;
}

// ---------------------------------------------------------------------------------------------
// Per-read driver (src/map.rs:56-111, 162-167; src/align.rs:34-55)
// ---------------------------------------------------------------------------------------------
struct ReadResult {
    u32 n_anchors = 0;
    std::vector<Chain> chains;
    AlnResult aln;
    u32 aln_chain = 0;
    std::vector<u64> path_handles;
    // debug
    std::vector<Anchor> anchors_emitted;
    ChainDebug cdbg;
    std::vector<u64> sub_handles;
    std::vector<std::pair<u32, u32>> sub_edges;
};

void map_one(const Index& ix, const char* read, u64 len, const orc_params_t& P, ReadResult& rr) {
    std::vector<Anchor> anchors;
    anchors_for_query(ix, read, len, !(P.flags & ORC_FLAG_ALL_ORIENTATIONS), anchors);   // src/map.rs:62 passes only_forward = true
    rr.n_anchors = (u32)anchors.size();
    const bool dbg = P.flags & ORC_FLAG_KEEP_DEBUG;
    if (dbg) rr.anchors_emitted = anchors;
    chain_anchors(anchors, ix.k, P.bandwidth, P.max_gap, P.min_anchors, rr.chains, dbg ? &rr.cdbg : nullptr);
    if (P.flags & ORC_FLAG_NO_ALIGN) return;
    // best_alignment_for_query (src/align.rs:34-55)
    size_t n_try = std::min<size_t>(P.best_n, rr.chains.size());
    bool have = false;
    for (size_t c = 0; c < n_try; c++) {
        const Chain& ch = rr.chains[c];
        if (ch.placeholder) continue;     // from_placeholder_chain: path_length None sorts last
        AlnResult a;
        std::vector<u64> handles;
        std::vector<std::string> labels;
        std::vector<std::pair<u32, u32>> edges;
        if (!find_range_chain(ix, ch, handles)) {
            a.status = ST_UNSUPPORTED;
        } else {
            extend_range(ix, ch, len, handles, P.flags & ORC_FLAG_LITERAL_EXTEND);
            nodes_edges(ix, handles, labels, edges);
            poa_align(labels, edges, read, (int32_t)len, P, a);
        }
        // stable sort by path_length descending, first wins (src/align.rs:52-54)
        if (!have || a.path_len > rr.aln.path_len) {
            rr.aln = a;
            rr.aln_chain = (u32)c;
            rr.path_handles.clear();
            for (u32 nidx : a.path_nodes) rr.path_handles.push_back(handles[nidx]);
            if (dbg) { rr.sub_handles = handles; rr.sub_edges = edges; }
            have = true;
        }
    }
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// C API
// ---------------------------------------------------------------------------------------------
struct orc_index { Index ix; };

struct orc_results {
    u32 n = 0;
    std::map<std::string, std::pair<const void*, u64>> fields;
    std::vector<u32> n_anchors, aln_chain, aln_path_len, aln_start_off, aln_end_off, aln_n_aligned;
    std::vector<u64> chain_off, chain_anchor_off, aln_path_off, aln_path, aln_run_off, aln_cells;
    std::vector<u32> ca_id, ca_qb, ca_tb, ca_te, aln_runs;
    std::vector<double> ca_score;
    std::vector<uint8_t> aln_status;
    std::vector<int32_t> aln_score;
    std::vector<u64> dbg_anchor_off, dbg_sub_off, dbg_sub_handles, dbg_edge_off;
    std::vector<u32> dbg_qb, dbg_tb, dbg_te, dbg_sorted, dbg_edges;
    std::vector<double> dbg_f;
    std::vector<int32_t> dbg_pred;
    std::string gaf[2];
    template <class T> void reg(const char* name, const std::vector<T>& v) { fields[name] = {v.data(), v.size()}; }
    void finish() {
        reg("n_anchors", n_anchors); reg("chain_off", chain_off); reg("chain_anchor_off", chain_anchor_off);
        reg("ca_id", ca_id); reg("ca_qb", ca_qb); reg("ca_tb", ca_tb); reg("ca_te", ca_te); reg("ca_score", ca_score);
        reg("aln_status", aln_status); reg("aln_chain", aln_chain); reg("aln_path_off", aln_path_off);
        reg("aln_path", aln_path); reg("aln_run_off", aln_run_off); reg("aln_runs", aln_runs);
        reg("aln_path_len", aln_path_len); reg("aln_start_off", aln_start_off); reg("aln_end_off", aln_end_off);
        reg("aln_n_aligned", aln_n_aligned); reg("aln_score", aln_score); reg("aln_cells", aln_cells);
        reg("dbg_anchor_off", dbg_anchor_off); reg("dbg_qb", dbg_qb); reg("dbg_tb", dbg_tb); reg("dbg_te", dbg_te);
        reg("dbg_sorted", dbg_sorted); reg("dbg_f", dbg_f); reg("dbg_pred", dbg_pred);
        reg("dbg_sub_off", dbg_sub_off); reg("dbg_sub_handles", dbg_sub_handles);
        reg("dbg_edge_off", dbg_edge_off); reg("dbg_edges", dbg_edges);
    }
    void append(const ReadResult& rr, bool dbg) {
        n_anchors.push_back(rr.n_anchors);
        for (const Chain& c : rr.chains) {
            if (c.placeholder) continue;     // a placeholder is represented by zero chains
            for (const Anchor& a : c.anchors) {
                ca_id.push_back((u32)a.id); ca_qb.push_back((u32)a.qb); ca_tb.push_back((u32)a.tb | (a.tb_rev ? 0x80000000u : 0u));
                ca_te.push_back((u32)a.te | (a.te_rev ? 0x80000000u : 0u)); ca_score.push_back(a.f);
            }
            chain_anchor_off.push_back(ca_id.size());
        }
        chain_off.push_back(chain_anchor_off.size() - 1);
        aln_status.push_back(rr.aln.status); aln_chain.push_back(rr.aln_chain);
        aln_path.insert(aln_path.end(), rr.path_handles.begin(), rr.path_handles.end());
        aln_path_off.push_back(aln_path.size());
        aln_runs.insert(aln_runs.end(), rr.aln.runs.begin(), rr.aln.runs.end());
        aln_run_off.push_back(aln_runs.size());
        aln_path_len.push_back(rr.aln.path_len); aln_start_off.push_back(rr.aln.start_off);
        aln_end_off.push_back(rr.aln.end_off); aln_n_aligned.push_back(rr.aln.n_aligned);
        aln_score.push_back(rr.aln.score); aln_cells.push_back(rr.aln.cells);
        if (dbg) {
            for (const Anchor& a : rr.anchors_emitted) {
                dbg_qb.push_back((u32)a.qb); dbg_tb.push_back((u32)a.tb | (a.tb_rev ? 0x80000000u : 0u));
                dbg_te.push_back((u32)a.te | (a.te_rev ? 0x80000000u : 0u));
            }
            dbg_sorted.insert(dbg_sorted.end(), rr.cdbg.sorted.begin(), rr.cdbg.sorted.end());
            dbg_f.insert(dbg_f.end(), rr.cdbg.f.begin(), rr.cdbg.f.end());
            dbg_pred.insert(dbg_pred.end(), rr.cdbg.pred.begin(), rr.cdbg.pred.end());
            dbg_anchor_off.push_back(dbg_qb.size());
            dbg_sub_handles.insert(dbg_sub_handles.end(), rr.sub_handles.begin(), rr.sub_handles.end());
            dbg_sub_off.push_back(dbg_sub_handles.size());
            for (auto& e : rr.sub_edges) { dbg_edges.push_back(e.first); dbg_edges.push_back(e.second); }
            dbg_edge_off.push_back(dbg_edges.size() / 2);
        }
    }
    void begin() {
        chain_off.assign(1, 0); chain_anchor_off.assign(1, 0); aln_path_off.assign(1, 0); aln_run_off.assign(1, 0);
        dbg_anchor_off.assign(1, 0); dbg_sub_off.assign(1, 0); dbg_edge_off.assign(1, 0);
    }
};

extern "C" {

void orc_default_params(orc_params_t* p) {
    p->bandwidth = 50; p->max_gap = 1000; p->min_anchors = 3; p->best_n = 1;
    p->match = 2; p->mismatch = 4; p->o1 = 4; p->e1 = 2; p->o2 = 24; p->e2 = 1;
    p->wb = 10; p->wf = 0.01f; p->flags = 0;
}

orc_index_t* orc_index_create(uint32_t k, uint64_t seq_len, uint64_t n_nodes, uint64_t n_edges,
                              const char* seq_fwd, const uint64_t* node_ref, const uint64_t* edges,
                              uint64_t n_keys, const char* keys, const uint64_t* key_offset, uint64_t n_tab,
                              const uint64_t* tab_start, const uint64_t* tab_end) {
    orc_index* o = new orc_index();
    Index& ix = o->ix;
    ix.k = k; ix.L = seq_len; ix.n_nodes = n_nodes; ix.n_edges = n_edges;
    ix.fwd.assign(seq_fwd, seq_len);
    ix.rev = revcomp(ix.fwd);
    for (u64 i = 0; i <= n_nodes; i++) {
        ix.seq_idx.push_back(node_ref[3 * i]); ix.edge_idx.push_back(node_ref[3 * i + 1]);
        ix.edges_in.push_back(node_ref[3 * i + 2]);
    }
    ix.edges.assign(edges, edges + n_edges);
    ix.key2off.reserve(n_keys * 2);
    for (u64 i = 0; i < n_keys; i++) ix.key2off.emplace(std::string(keys + i * k, k), key_offset[i]);
    ix.tab_start.assign(tab_start, tab_start + n_tab);
    ix.tab_end.assign(tab_end, tab_end + n_tab);
    return o;
}
void orc_index_free(orc_index_t* o) { delete o; }

uint64_t orc_bv_rank(const orc_index_t* o, uint64_t pos) { return o->ix.rank(pos); }
uint64_t orc_bv_inverse_rank(const orc_index_t* o, uint64_t pos) { return o->ix.inverse_rank(pos); }
uint64_t orc_bv_select(const orc_index_t* o, uint64_t n) { return o->ix.select(n); }
uint64_t orc_handle_from_seqpos(const orc_index_t* o, int reverse, uint64_t pos) {
    return o->ix.handle_from_seqpos(reverse != 0, pos);
}
static uint64_t copy_out(const std::vector<u64>& v, uint64_t* out, uint64_t cap) {
    for (size_t i = 0; i < v.size() && i < cap; i++) out[i] = v[i];
    return v.size();
}
uint64_t orc_incoming(const orc_index_t* o, uint64_t h, uint64_t* out, uint64_t cap) { return copy_out(o->ix.incoming(h), out, cap); }
uint64_t orc_outgoing(const orc_index_t* o, uint64_t h, uint64_t* out, uint64_t cap) { return copy_out(o->ix.outgoing(h), out, cap); }
uint64_t orc_seq_from_handle(const orc_index_t* o, uint64_t h, char* out, uint64_t cap) {
    std::string s = o->ix.seq_from_handle(h);
    memcpy(out, s.data(), std::min<u64>(cap, s.size()));
    return s.size();
}
uint64_t orc_positions(const orc_index_t* o, const char* kmer, uint64_t len, uint64_t* start, uint64_t* end, uint64_t cap) {
    std::vector<u64> s, e;
    positions_for(o->ix, kmer, len, s, e);
    copy_out(s, start, cap); copy_out(e, end, cap);
    return s.size();
}
uint64_t orc_anchors(const orc_index_t* o, const char* read, uint64_t len, int only_forward, uint64_t* qb, uint64_t* qe,
                     uint64_t* tb, uint64_t* te, uint64_t cap) {
    std::vector<Anchor> a;
    anchors_for_query(o->ix, read, len, only_forward != 0, a);
    for (size_t i = 0; i < a.size() && i < cap; i++) {
        qb[i] = a[i].qb; qe[i] = a[i].qe;
        tb[i] = a[i].tb | (a[i].tb_rev ? REV_BIT : 0);
        te[i] = a[i].te | (a[i].te_rev ? REV_BIT : 0);
    }
    return a.size();
}
double orc_score_anchor(uint64_t a_qb, uint64_t a_qe, uint64_t a_tb, uint64_t a_te, double a_f, uint64_t b_qb, uint64_t b_qe,
                        uint64_t b_tb, uint64_t b_te, uint64_t seed_length, uint64_t max_gap) {
    Anchor a{0, a_qb, a_qe, a_tb & ~REV_BIT, a_te & ~REV_BIT, (a_tb & REV_BIT) != 0, (a_te & REV_BIT) != 0, a_f, -1};
    Anchor b{1, b_qb, b_qe, b_tb & ~REV_BIT, b_te & ~REV_BIT, (b_tb & REV_BIT) != 0, (b_te & REV_BIT) != 0, 0.0, -1};
    return score_anchor(a, b, seed_length, max_gap);
}

orc_results_t* orc_map_reads(const orc_index_t* o, const char* seqs, const uint64_t* seq_off, uint32_t n_reads,
                             const orc_params_t* P, int n_threads) {
    orc_results* R = new orc_results();
    R->n = n_reads;
    R->begin();
    const bool dbg = P->flags & ORC_FLAG_KEEP_DEBUG;
    const u32 CH = 4096;
    std::vector<ReadResult> buf;
    for (u32 b = 0; b < n_reads; b += CH) {
        u32 e = std::min(n_reads, b + CH);
        buf.assign(e - b, ReadResult());
        // reads are independent: worker threads pull blocks of 16 reads from a shared counter
        std::atomic<u32> next(b);
        auto worker = [&]() {
            for (;;) {
                u32 r0 = next.fetch_add(16);
                if (r0 >= e) break;
                for (u32 r = r0; r < std::min(e, r0 + 16); r++)
                    map_one(o->ix, seqs + seq_off[r], seq_off[r + 1] - seq_off[r], *P, buf[r - b]);
            }
        };
        const int nt = n_threads > 1 ? n_threads : 1;
        std::vector<std::thread> pool;
        for (int t = 1; t < nt; t++) pool.emplace_back(worker);
        worker();
        for (auto& t : pool) t.join();
        for (u32 r = b; r < e; r++) R->append(buf[r - b], dbg);
    }
    R->finish();
    return R;
}

orc_results_t* orc_poa_align(const char* labels, const uint64_t* label_off, uint32_t n_nodes, const uint32_t* edges,
                             uint32_t n_edges, const char* query, uint32_t qlen, const orc_params_t* P) {
    std::vector<std::string> lab;
    for (u32 i = 0; i < n_nodes; i++) lab.emplace_back(labels + label_off[i], label_off[i + 1] - label_off[i]);
    std::vector<std::pair<u32, u32>> ed;
    for (u32 i = 0; i < n_edges; i++) ed.push_back({edges[2 * i], edges[2 * i + 1]});
    ReadResult rr;
    poa_align(lab, ed, query, (int32_t)qlen, *P, rr.aln);
    for (u32 nidx : rr.aln.path_nodes) rr.path_handles.push_back(nidx);   // node indices stand for handles
    orc_results* R = new orc_results();
    R->n = 1;
    R->begin();
    R->append(rr, false);
    R->finish();
    return R;
}

void orc_results_free(orc_results_t* r) { delete r; }

const void* orc_results_field(const orc_results_t* r, const char* name, uint64_t* count) {
    auto it = r->fields.find(name);
    if (it == r->fields.end()) { if (count) *count = 0; return nullptr; }
    if (count) *count = it->second.second;
    return it->second.first;
}

// GAF text: GAFAlignment::{from_chain, from_placeholder_chain, to_string} (src/align.rs:762-1027) and
// generate_alignment (src/align.rs:1096-1168).
const char* orc_results_gaf(orc_results_t* R, const orc_index_t* o, const char* names, const char* seqs,
                            const uint64_t* seq_off, int which, uint64_t* len) {
    const Index& ix = o->ix;
    std::vector<std::string> nm;
    {
        std::string all(names);
        size_t b = 0;
        for (u32 i = 0; i < R->n; i++) {
            size_t e = all.find('\n', b);
            nm.push_back(all.substr(b, e == std::string::npos ? std::string::npos : e - b));
            b = e == std::string::npos ? all.size() : e + 1;
        }
    }
    const bool cs_prefix = !(which & 2);          // which: bit 0 = alignments file, bit 1 = cs string without its "cs:Z:" tag (ORC_FLAG_NO_CS_PREFIX)
    which &= 1;
    std::string& out = R->gaf[which ? 1 : 0];
    out.clear();
    auto placeholder = [&](u32 r) {
        u64 l = seq_off[r + 1] - seq_off[r];
        out += nm[r] + "\t" + std::to_string(l) + "\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n";       // :913-930
    };
    for (u32 r = 0; r < R->n; r++) {
        u64 qlen = seq_off[r + 1] - seq_off[r];
        if (which == 0) {
            if (R->chain_off[r] == R->chain_off[r + 1]) { placeholder(r); continue; }
            for (u64 c = R->chain_off[r]; c < R->chain_off[r + 1]; c++) {
                u64 b = R->chain_anchor_off[c], e = R->chain_anchor_off[c + 1];
                std::string path;
                for (u64 a = b; a < e; a++) {        // AnchorPosOnGraph::new (src/chain.rs:89-127)
                    // bit 31 = Reverse.  AnchorPosOnGraph::new subtracts the FORWARD start of the node (get_bv_select) also from a
                    // reverse-strand position (src/chain.rs:96-121): u64 arithmetic, wrapping as a release build does
                    const bool r1 = R->ca_tb[a] >> 31, r2 = R->ca_te[a] >> 31;
                    u64 tb = R->ca_tb[a] & 0x7fffffffu, tei = (u64)(R->ca_te[a] & 0x7fffffffu) - 1;
                    u64 h1 = ix.handle_from_seqpos(r1, tb), h2 = ix.handle_from_seqpos(r2, tei);
                    path += std::string("(") + (r1 ? "<" : ">") + std::to_string(h1 >> 1) + ":" + std::to_string(tb - ix.select(h1 >> 1)) + "," +
                            (r2 ? "<" : ">") + std::to_string(h2 >> 1) + ":" + std::to_string(tei - ix.select(h2 >> 1)) + "),";
                }
                out += nm[r] + "\t" + std::to_string(qlen) + "\t" + std::to_string(R->ca_qb[b]) + "\t" +
                       std::to_string(R->ca_qb[e - 1] + ix.k) + "\t+\t" + path + "\t0\t0\t0\t0\t0\t0\t" +
                       "ta:Z:chain,n_anchors: " + std::to_string(e - b) + "\n";                  // :892-910
            }
        } else {
            uint8_t st = R->aln_status[r];
            if (st == ST_PLACEHOLDER) { placeholder(r); continue; }
            std::string path, cs = cs_prefix ? "cs:Z:" : "", cg;
            for (u64 p = R->aln_path_off[r]; p < R->aln_path_off[r + 1]; p++) {
                u64 h = R->aln_path[p];
                path += ((h & 1) ? "<" : ">") + std::to_string(h >> 1);                          // :1127-1133
            }
            // walk the runs to rebuild cs / cigar
            const char* q = seqs + seq_off[r];
            u64 pi = R->aln_path_off[r], qi = 0;
            std::string node_seq = pi < R->aln_path_off[r + 1] ? ix.seq_from_handle(R->aln_path[pi]) : std::string();
            u64 off = R->aln_start_off[r];
            auto next_base = [&]() -> char {
                while (off >= node_seq.size()) { pi++; node_seq = ix.seq_from_handle(R->aln_path[pi]); off = 0; }
                return node_seq[off++];
            };
            u32 mrun = 0;
            auto flush_m = [&]() { if (mrun) { cg += std::to_string(mrun) + "M"; mrun = 0; } };
            for (u64 t = R->aln_run_off[r]; t < R->aln_run_off[r + 1]; t++) {
                u32 op = R->aln_runs[t] >> 28, n = R->aln_runs[t] & 0x0fffffffu;
                if (op == 0) { cs += ":" + std::to_string(n); for (u32 x = 0; x < n; x++) { next_base(); qi++; } mrun += n; }
                else if (op == 1) { for (u32 x = 0; x < n; x++) { cs += "*"; cs += (char)tolower(next_base()); cs += (char)tolower(q[qi++]); } mrun += n; }
                else if (op == 2) { flush_m(); cs += "-"; for (u32 x = 0; x < n; x++) cs += (char)tolower(next_base()); cg += std::to_string(n) + "D"; }
                else { flush_m(); cs += "+"; for (u32 x = 0; x < n; x++) cs += (char)tolower(q[qi++]); cg += std::to_string(n) + "I"; }
            }
            flush_m();
            out += nm[r] + "\t" + std::to_string(qlen) + "\t0\t" + std::to_string(qlen) + "\t+\t" + path + "\t" +
                   std::to_string(R->aln_path_len[r]) + "\t" + std::to_string(R->aln_start_off[r]) + "\t" +
                   std::to_string(R->aln_end_off[r]) + "\t0\t" + std::to_string(R->aln_n_aligned[r]) + "\t255\t" +
                   "as:i:-30 " + cs + ",cg:Z:" + cg + "\n";                                     // :1145-1167
        }
    }
    if (len) *len = out.size();
    return out.c_str();
}

}  // extern "C"
