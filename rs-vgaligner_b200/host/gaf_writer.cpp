// GAF text from the C ABI's flat result arrays, formatted on all host cores.
//
// Records follow the reference byte for byte:
//   chains file      GAFAlignment::from_chain (src/align.rs:762-911): name, qlen, first anchor's query_begin, last
//                    anchor's query_end, '+', one "(>node:offset,>node:offset)," group per anchor (start and INCLUSIVE end
//                    position of the anchor on the graph, src/chain.rs:89-127), six zero columns, "ta:Z:chain,n_anchors: N"
//   alignments file  generate_alignment (src/align.rs:1096-1168): name, qlen, 0, qlen, '+', >id/<id path, |abpoa_nodes|,
//                    aln_start_offset, aln_end_offset, 0, n_aligned_bases, 255, "as:i:-30 " + cs + ",cg:Z:" + cigar
//   placeholder      GAFAlignment::from_placeholder_chain (src/align.rs:913-930): name, qlen, nine '*', 0, '*'
//   line layout      GAFAlignment::to_string (src/align.rs:971-1027): 13 tab-separated columns, '\n'
// A read whose device alignment did not complete (aln_status other than OK / PLACEHOLDER: band error, band wider than
// the widest kernel, ...) has no counterpart among the reference's records; it is written as a placeholder line and
// counted in *n_failed so that the caller can report it (never a well-formed alignment record with empty fields).
//
// Work is cut into contiguous read ranges of equal output volume (anchors for the chains file, runs + path steps for
// the alignments file); every thread formats its range into its own buffer with hand-rolled integer formatting (no
// std::string temporaries), and the buffers are copied into one allocation in parallel.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "flat_index.h"
#include "vga_host.h"

namespace vga {

// position -> node id (rank over seq_bv, src/index.rs:427-439) for every forward position; built once per index
struct GafTables {
    std::vector<uint32_t> pos2node;
    std::vector<uint64_t> starts;     // node id - 1 -> first position (select, src/index.rs:462-480)
};

static void build_tables(const FlatIndex& ix, GafTables& t) {
    t.starts.resize(ix.n_nodes + 1);
    for (uint64_t i = 0; i <= ix.n_nodes; i++) t.starts[i] = ix.seq_idx(i);
    t.pos2node.resize(ix.seq_len);
    for (uint64_t id = 1; id <= ix.n_nodes; id++)
        for (uint64_t p = t.starts[id - 1]; p < t.starts[id]; p++) t.pos2node[p] = (uint32_t)id;
}

struct Buf {
    std::vector<char> v;
    size_t n = 0;
    void need(size_t extra) {
        if (n + extra > v.size()) v.resize(std::max(v.size() * 2, n + extra + (1u << 16)));
    }
    char* cur() { return v.data() + n; }
};

// buffers of finished calls, reused by later ones (any thread); at most POOL_MAX of at most POOL_BYTES each are kept
static std::mutex g_buf_mu;
static std::vector<Buf*> g_buf_pool;
static const size_t POOL_MAX = 64, POOL_BYTES = 256u << 20;
struct BufLease {
    std::vector<Buf*> b;
    explicit BufLease(int n) : b((size_t)n, nullptr) {
        std::lock_guard<std::mutex> lk(g_buf_mu);
        for (auto& x : b) {
            if (!g_buf_pool.empty()) { x = g_buf_pool.back(); g_buf_pool.pop_back(); }
            else x = new Buf();
        }
    }
    ~BufLease() {
        std::lock_guard<std::mutex> lk(g_buf_mu);
        for (Buf* x : b) {
            if (g_buf_pool.size() < POOL_MAX && x->v.size() <= POOL_BYTES) g_buf_pool.push_back(x);
            else delete x;
        }
    }
    Buf& operator[](int t) { return *b[(size_t)t]; }
};

static const char DIG2[201] =
    "0001020304050607080910111213141516171819202122232425262728293031323334353637383940414243444546474849"
    "5051525354555657585960616263646566676869707172737475767778798081828384858687888990919293949596979899";

static inline char* put_u64(char* p, uint64_t v) {
    // the common sizes (offsets inside a node, read positions, node ids) straight into place, two digits at a time
    if (v < 100) {
        if (v < 10) { *p = (char)('0' + v); return p + 1; }
        memcpy(p, DIG2 + 2 * v, 2);
        return p + 2;
    }
    if (v < 10000) {
        const unsigned hi = (unsigned)v / 100, lo = (unsigned)v % 100;
        if (hi < 10) { *p++ = (char)('0' + hi); } else { memcpy(p, DIG2 + 2 * hi, 2); p += 2; }
        memcpy(p, DIG2 + 2 * lo, 2);
        return p + 2;
    }
    if (v < 100000000ull) {
        const unsigned a = (unsigned)(v / 10000), b = (unsigned)(v % 10000);
        p = put_u64(p, a);                                   // 1 .. 4 digits
        memcpy(p, DIG2 + 2 * (b / 100), 2);
        memcpy(p + 2, DIG2 + 2 * (b % 100), 2);
        return p + 4;
    }
    char tmp[20];
    int n = 0;
    while (v >= 100) { const uint64_t q = v / 100; const unsigned r = (unsigned)(v - q * 100); tmp[n++] = DIG2[2 * r + 1]; tmp[n++] = DIG2[2 * r]; v = q; }
    if (v >= 10) { tmp[n++] = DIG2[2 * v + 1]; tmp[n++] = DIG2[2 * v]; } else tmp[n++] = (char)('0' + v);
    while (n) *p++ = tmp[--n];
    return p;
}
static inline char* put_str(char* p, const char* s, size_t n) { memcpy(p, s, n); return p + n; }
#define PUT_LIT(p, lit) put_str(p, lit, sizeof(lit) - 1)

struct Names {
    const char* joined = nullptr;                 // '\n' separated, or null: "read<base+i>"
    std::vector<const char*> beg;                 // start of every name (+ one past the last)
    uint64_t base = 0;
    void index(uint32_t n) {
        if (!joined) return;
        beg.reserve((size_t)n + 1);
        const char* p = joined;
        for (uint32_t i = 0; i < n; i++) {
            beg.push_back(p);
            const char* e = strchr(p, '\n');
            p = e ? e + 1 : p + strlen(p) + 1;
        }
        beg.push_back(p);
    }
    size_t len(uint32_t i) const { return joined ? (size_t)(beg[i + 1] - beg[i] - 1) : 24; }
    char* put(char* p, uint32_t i) const {
        if (joined) return put_str(p, beg[i], (size_t)(beg[i + 1] - beg[i] - 1));
        p = PUT_LIT(p, "read");
        return put_u64(p, base + i);
    }
};

static inline char lower(char c) { return (c >= 'A' && c <= 'Z') ? (char)(c + 32) : c; }

struct Job {
    const FlatIndex* ix;
    const GafTables* tb;
    const VgaResultsView* R;
    const Names* names;
    const char* seqs;
    const uint64_t* off;
    int which;
    bool cs_prefix;   // alignments file: the cs string with its "cs:Z:" tag (rs-abPOA's exact format is [RECALLED]: a switch, like the oracle's)
};

static inline char* put_placeholder(char* p, const Job& J, uint32_t r) {
    p = J.names->put(p, r);
    *p++ = '\t';
    p = put_u64(p, J.off[r + 1] - J.off[r]);
    return PUT_LIT(p, "\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n");
}

static void format_chains(const Job& J, uint32_t r0, uint32_t r1, Buf& out) {
    const VgaResultsView& R = *J.R;
    const uint32_t k = J.ix->k;
    uint64_t c_lo = 1, c_hi = 0;              // forward positions [c_lo, c_hi) of the cached node (empty at first)
    uint32_t c_node = 0;
    char c_txt[24];
    size_t c_len = 0;
    for (uint32_t r = r0; r < r1; r++) {
        const size_t nl = J.names->len(r);
        if (R.chain_off[r] == R.chain_off[r + 1]) {
            out.need(nl + 64);
            out.n = (size_t)(put_placeholder(out.cur(), J, r) - out.v.data());
            continue;
        }
        for (uint64_t c = R.chain_off[r]; c < R.chain_off[r + 1]; c++) {
            const uint64_t b = R.chain_anchor_off[c], e = R.chain_anchor_off[c + 1];
            out.need(nl + 160 + (e - b) * 92);        // "(>" 20 ":" 20 ",>" 20 ":" 20 ")," per anchor at the very most
            char* p = out.cur();
            p = J.names->put(p, r);
            *p++ = '\t';
            p = put_u64(p, J.off[r + 1] - J.off[r]);
            *p++ = '\t';
            p = put_u64(p, R.ca_qb[b]);
            *p++ = '\t';
            p = put_u64(p, (uint64_t)R.ca_qb[e - 1] + k);
            p = PUT_LIT(p, "\t+\t");
            for (uint64_t a = b; a < e; a++) {
                // bit 31 = Reverse (only with VGA_FLAG_ALL_ORIENTATIONS).  The node of a reverse-strand position p is the node of the
                // forward position L - 1 - p (handle_from_seqpos, src/index.rs:403-405); AnchorPosOnGraph::new subtracts the node's
                // FORWARD start also from a reverse-strand position (src/chain.rs:96-121): u64 arithmetic, wrapping like a release build
                const bool r1 = R.ca_tb[a] >> 31, r2 = R.ca_te[a] >> 31;
                const uint64_t tb = R.ca_tb[a] & 0x7fffffffu, tei = (uint64_t)(R.ca_te[a] & 0x7fffffffu) - 1;      // get_end_seqpos_inclusive, src/chain.rs:65-70
                // consecutive anchors of a chain mostly stay on one node: the node of the previous position is tried first, and its
                // decimal id is kept as text
                const uint64_t f1 = r1 ? J.ix->seq_len - 1 - tb : tb, f2 = r2 ? J.ix->seq_len - 1 - tei : tei;
                if (f1 < c_lo || f1 >= c_hi) {
                    c_node = J.tb->pos2node[f1];
                    c_lo = J.tb->starts[c_node - 1]; c_hi = J.tb->starts[c_node];
                    c_len = (size_t)(put_u64(c_txt, c_node) - c_txt);
                }
                const uint64_t s1 = c_lo;
                *p++ = '('; *p++ = r1 ? '<' : '>';
                memcpy(p, c_txt, 16); p += c_len;                  // a u32 id has at most 10 digits; the 16 bytes are inside the reserve
                *p++ = ':';
                p = put_u64(p, tb - s1);
                if (f2 < c_lo || f2 >= c_hi) {
                    c_node = J.tb->pos2node[f2];
                    c_lo = J.tb->starts[c_node - 1]; c_hi = J.tb->starts[c_node];
                    c_len = (size_t)(put_u64(c_txt, c_node) - c_txt);
                }
                *p++ = ','; *p++ = r2 ? '<' : '>';
                memcpy(p, c_txt, 16); p += c_len;
                *p++ = ':';
                p = put_u64(p, tei - c_lo);
                *p++ = ')'; *p++ = ',';
            }
            p = PUT_LIT(p, "\t0\t0\t0\t0\t0\t0\tta:Z:chain,n_anchors: ");
            p = put_u64(p, e - b);
            *p++ = '\n';
            out.n = (size_t)(p - out.v.data());
        }
    }
}

static void format_alignments(const Job& J, uint32_t r0, uint32_t r1, Buf& out, uint64_t& failed) {
    const VgaResultsView& R = *J.R;
    const FlatIndex& ix = *J.ix;
    std::vector<char> cg;
    for (uint32_t r = r0; r < r1; r++) {
        const size_t nl = J.names->len(r);
        const uint64_t qlen = J.off[r + 1] - J.off[r];
        if (R.aln_status[r] != VGA_ALN_OK) {
            if (R.aln_status[r] != VGA_ALN_PLACEHOLDER) failed++;
            out.need(nl + 64);
            out.n = (size_t)(put_placeholder(out.cur(), J, r) - out.v.data());
            continue;
        }
        const uint64_t p0 = R.aln_path_off[r], p1 = R.aln_path_off[r + 1], t0 = R.aln_run_off[r], t1 = R.aln_run_off[r + 1];
        // cs: <= 3 bytes per mismatch column, 1 per deleted / inserted base, 11 per run; cigar: 11 per run
        out.need(nl + 256 + (p1 - p0) * 22 + 3 * (qlen + R.aln_path_len[r]) + (t1 - t0) * 24);
        char* p = out.cur();
        p = J.names->put(p, r);
        *p++ = '\t';
        p = put_u64(p, qlen);
        p = PUT_LIT(p, "\t0\t");
        p = put_u64(p, qlen);
        p = PUT_LIT(p, "\t+\t");
        for (uint64_t x = p0; x < p1; x++) { *p++ = (R.aln_path[x] & 1) ? '<' : '>'; p = put_u64(p, R.aln_path[x] >> 1); }
        *p++ = '\t';
        p = put_u64(p, R.aln_path_len[r]);
        *p++ = '\t';
        p = put_u64(p, R.aln_start_off[r]);
        *p++ = '\t';
        p = put_u64(p, R.aln_end_off[r]);
        p = PUT_LIT(p, "\t0\t");
        p = put_u64(p, R.aln_n_aligned[r]);
        p = PUT_LIT(p, "\t255\tas:i:-30 ");
        if (J.cs_prefix) p = PUT_LIT(p, "cs:Z:");
        // walk the path's bases: label of a forward handle is a slice of seq_fwd, of a reverse handle the mirrored slice
        // of seq_rev (seq_from_handle, src/index.rs:503-533)
        const char* q = J.seqs + J.off[r];
        uint64_t pi = p0, qi = 0;
        const char* nseq = nullptr;
        uint64_t nlen = 0, noff = R.aln_start_off[r];
        auto load = [&](uint64_t hd) {
            const uint64_t id = hd >> 1, b = ix.seq_idx(id - 1), e = ix.seq_idx(id);
            nseq = (hd & 1) ? ix.seq_rev.data() + (ix.seq_len - e) : ix.seq_fwd.data() + b;
            nlen = e - b;
        };
        if (pi < p1) load(R.aln_path[pi]);
        auto next_base = [&]() -> char {
            while (noff >= nlen) { pi++; load(R.aln_path[pi]); noff = 0; }
            return nseq[noff++];
        };
        cg.clear();
        char num[24];
        uint32_t mrun = 0;
        auto cg_put = [&](uint32_t n, char op) { char* e = put_u64(num, n); cg.insert(cg.end(), num, e); cg.push_back(op); };
        for (uint64_t t = t0; t < t1; t++) {
            const uint32_t op = R.aln_runs[t] >> 28, n = R.aln_runs[t] & 0x0fffffffu;
            if (op == 0) {
                *p++ = ':';
                p = put_u64(p, n);
                for (uint32_t x = 0; x < n; x++) next_base();
                qi += n; mrun += n;
            } else if (op == 1) {
                for (uint32_t x = 0; x < n; x++) { *p++ = '*'; *p++ = lower(next_base()); *p++ = lower(q[qi++]); }
                mrun += n;
            } else if (op == 2) {
                if (mrun) { cg_put(mrun, 'M'); mrun = 0; }
                *p++ = '-';
                for (uint32_t x = 0; x < n; x++) *p++ = lower(next_base());
                cg_put(n, 'D');
            } else {
                if (mrun) { cg_put(mrun, 'M'); mrun = 0; }
                *p++ = '+';
                for (uint32_t x = 0; x < n; x++) *p++ = lower(q[qi++]);
                cg_put(n, 'I');
            }
        }
        if (mrun) cg_put(mrun, 'M');
        p = PUT_LIT(p, ",cg:Z:");
        p = put_str(p, cg.data(), cg.size());
        *p++ = '\n';
        out.n = (size_t)(p - out.v.data());
    }
}

char* gaf_format(const FlatIndex& ix, const GafTables& tb, const VgaResultsView* R, const char* names_joined, uint64_t name_base,
                 const char* seqs, const uint64_t* off, int which, int threads, uint64_t* len, uint64_t* n_failed) {
    const uint32_t n = R->n_reads;
    Names names;
    names.joined = names_joined; names.base = name_base;
    names.index(n);
    const bool cs_prefix = !(which & 2);      // which: bit 0 = alignments file, bit 1 = no "cs:Z:" tag
    which &= 1;
    Job J{&ix, &tb, R, &names, seqs, off, which, cs_prefix};
    int T = threads > 0 ? threads : (int)std::max(1u, std::thread::hardware_concurrency());
    // ranges of equal output volume
    auto weight = [&](uint32_t r) -> uint64_t {
        if (which == 0) {
            const uint64_t c0 = R->chain_off[r], c1 = R->chain_off[r + 1];
            return 4 + (c1 > c0 ? R->chain_anchor_off[c1] - R->chain_anchor_off[c0] : 0);
        }
        return 8 + (R->aln_run_off[r + 1] - R->aln_run_off[r]) + (R->aln_path_off[r + 1] - R->aln_path_off[r]) + (off[r + 1] - off[r]) / 16;
    };
    uint64_t total = 0;
    for (uint32_t r = 0; r < n; r++) total += weight(r);
    if (total < (1u << 16)) T = 1;
    T = (int)std::min<uint64_t>((uint64_t)T, std::max<uint32_t>(n, 1));
    std::vector<uint32_t> cut(T + 1, n);
    cut[0] = 0;
    {
        uint64_t acc = 0;
        int t = 1;
        for (uint32_t r = 0; r < n && t < T; r++) {
            acc += weight(r);
            while (t < T && acc >= total * (uint64_t)t / (uint64_t)T) cut[t++] = r + 1;
        }
    }
    // per-thread buffers come from a small pool and keep their size between calls: a fresh buffer is zero-filled and page-faulted
    // in by 16 threads at once, which cost more than the formatting itself
    BufLease bufs(T);
    std::vector<uint64_t> failed(T, 0);
    auto work = [&](int t) {
        const size_t want = (size_t)((which == 0 ? 40 : 24) * (total / (uint64_t)T + 64)) + (1u << 16);
        bufs[t].n = 0;
        if (bufs[t].v.size() < want) bufs[t].v.resize(want);
        if (which == 0) format_chains(J, cut[t], cut[t + 1], bufs[t]);
        else format_alignments(J, cut[t], cut[t + 1], bufs[t], failed[t]);
    };
    if (T == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    size_t tot = 0;
    std::vector<size_t> at(T);
    for (int t = 0; t < T; t++) { at[t] = tot; tot += bufs[t].n; }
    char* c = (char*)malloc(tot + 1);
    if (!c) return nullptr;
    auto copy = [&](int t) { memcpy(c + at[t], bufs[t].v.data(), bufs[t].n); };
    if (T == 1) copy(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back(copy, t);
        for (auto& x : th) x.join();
    }
    c[tot] = 0;
    if (len) *len = tot;
    if (n_failed) { *n_failed = 0; for (uint64_t f : failed) *n_failed += f; }
    return c;
}

const GafTables* gaf_tables_new(const FlatIndex& ix) {
    GafTables* t = new GafTables();
    build_tables(ix, *t);
    return t;
}
void gaf_tables_free(const GafTables* t) { delete t; }

}  // namespace vga
