// Host-side restatement of `vgaligner index` (GFA -> HashGraph -> Index), producing the flat
// arrays the device path consumes.  Not on the GPU hot path: the Rust front end keeps this job in
// the real deployment; this file exists because no `.idx` can be produced without cargo here.
//
// Follows (file:line under the reference tree):
//   HashGraph::from_gfa / create_edge / handle_edges_iter   handlegraph 0.5.0 (unvendored; restated
//                                                           from its published source, pinned by the
//                                                           reference tests at src/index.rs:1260-1367)
//   find_forward_sequence          src/utils.rs:81-146
//   reverse_complement             src/dna.rs:5-40
//   generate_kmers_parallel        src/kmer.rs:277-304
//   generate_kmer_with_handle_orient   src/kmer.rs:347-505
//   get_seq_pos / generate_pos_on_ref_2    src/kmer.rs:752-770, 816-928
#include "flat_index.h"

#include <algorithm>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace vga {

// ---------------------------------------------------------------------------------------------
// DNA helpers
// ---------------------------------------------------------------------------------------------

// vgaligner's own complement: acgtu/ACGTU complemented, every other byte -> 'N'
// (src/dna.rs:19-33); bytes outside [ACGTUacgtuN] are rejected (src/dna.rs:10-13 panics).
std::string reverse_complement(const std::string& s) {
    std::string out;
    out.reserve(s.size());
    for (size_t i = s.size(); i-- > 0;) {
        char c = s[i], r;
        switch (c) {
            case 'a': r = 't'; break;
            case 'c': r = 'g'; break;
            case 't': r = 'a'; break;
            case 'g': r = 'c'; break;
            case 'u': r = 'a'; break;
            case 'A': r = 'T'; break;
            case 'C': r = 'G'; break;
            case 'T': r = 'A'; break;
            case 'G': r = 'C'; break;
            case 'U': r = 'A'; break;
            case 'N': r = 'N'; break;
            default:
                throw std::runtime_error(std::string("Input sequence base is not DNA: ") + c);
        }
        out.push_back(r);
    }
    return out;
}

// handlegraph's complement used by HashGraph::sequence() on reverse handles: ACGT/acgt swapped,
// everything else (N, IUPAC) left as is.
static inline char hg_comp(char c) {
    switch (c) {
        case 'A': return 'T';
        case 'C': return 'G';
        case 'G': return 'C';
        case 'T': return 'A';
        case 'a': return 't';
        case 'c': return 'g';
        case 'g': return 'c';
        case 't': return 'a';
        default: return c;
    }
}

bool pack_kmer(const char* s, uint32_t k, uint64_t* out) {
    uint64_t v = 0;
    for (uint32_t i = 0; i < k; i++) {
        uint64_t c;
        switch (s[i]) {
            case 'A': c = 0; break;
            case 'C': c = 1; break;
            case 'G': c = 2; break;
            case 'T': c = 3; break;
            default: return false;
        }
        v = (v << 2) | c;
    }
    *out = v;
    return true;
}

// ---------------------------------------------------------------------------------------------
// HashGraph
// ---------------------------------------------------------------------------------------------

std::string HashGraph::sequence(handle_t h) const {
    const std::string& s = seq[handle_id(h)];
    if (!handle_rev(h)) return s;
    std::string out(s.size(), 'N');
    for (size_t i = 0; i < s.size(); i++) out[i] = hg_comp(s[s.size() - 1 - i]);
    return out;
}

std::vector<handle_t> HashGraph::edges_right(handle_t h) const {
    if (!handle_rev(h)) return right[handle_id(h)];
    std::vector<handle_t> out = left[handle_id(h)];
    for (auto& x : out) x = handle_flip(x);
    return out;
}

std::vector<handle_t> HashGraph::edges_left(handle_t h) const {
    if (!handle_rev(h)) return left[handle_id(h)];
    std::vector<handle_t> out = right[handle_id(h)];
    for (auto& x : out) x = handle_flip(x);
    return out;
}

// handlegraph 0.5.0 HashGraph::create_edge: the duplicate test looks at the left node's
// right_edges whatever the orientation of `l`; the mirrored entry is skipped for a
// self-inverting edge.
void HashGraph::create_edge(handle_t l, handle_t r) {
    const auto& lr = right[handle_id(l)];
    if (std::find(lr.begin(), lr.end(), r) != lr.end()) return;
    if (handle_rev(l)) left[handle_id(l)].push_back(handle_flip(r));
    else right[handle_id(l)].push_back(r);
    if (l != handle_flip(r)) {
        if (handle_rev(r)) right[handle_id(r)].push_back(handle_flip(l));
        else left[handle_id(r)].push_back(l);
    }
}

static std::vector<std::string> split_tabs(const std::string& line) {
    std::vector<std::string> f;
    size_t b = 0;
    while (true) {
        size_t e = line.find('\t', b);
        if (e == std::string::npos) { f.push_back(line.substr(b)); break; }
        f.push_back(line.substr(b, e - b));
        b = e + 1;
    }
    return f;
}

static uint64_t parse_id(const std::string& s) {
    if (s.empty()) throw std::runtime_error("GFA: empty segment name");
    uint64_t v = 0;
    for (char c : s) {
        if (c < '0' || c > '9') throw std::runtime_error("GFA: segment name is not an integer: " + s);
        v = v * 10 + (uint64_t)(c - '0');
    }
    return v;
}

HashGraph graph_from_gfa_text(const std::string& text) {
    struct Seg { uint64_t id; std::string seq; };
    struct Link { uint64_t from; bool frev; uint64_t to; bool trev; };
    std::vector<Seg> segs;
    std::vector<Link> links;
    std::istringstream in(text);
    std::string line;
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) continue;
        if (line[0] == 'S') {
            auto f = split_tabs(line);
            if (f.size() < 3) throw std::runtime_error("GFA: malformed S line");
            segs.push_back({parse_id(f[1]), f[2]});
        } else if (line[0] == 'L') {
            auto f = split_tabs(line);
            if (f.size() < 5) throw std::runtime_error("GFA: malformed L line");
            if ((f[2] != "+" && f[2] != "-") || (f[4] != "+" && f[4] != "-"))
                throw std::runtime_error("GFA: malformed orientation in L line");
            links.push_back({parse_id(f[1]), f[2] == "-", parse_id(f[3]), f[4] == "-"});
        }
    }
    HashGraph g;
    uint64_t max_id = 0;
    for (auto& s : segs) max_id = std::max(max_id, s.id);
    if (max_id != segs.size())
        throw std::runtime_error("GFA: segment ids must be contiguous 1..N (README.md:24-28)");
    g.n_nodes = max_id;
    g.seq.assign(max_id + 1, std::string());
    g.left.assign(max_id + 1, {});
    g.right.assign(max_id + 1, {});
    std::vector<char> seen(max_id + 1, 0);
    for (auto& s : segs) {
        if (s.id == 0 || seen[s.id]) throw std::runtime_error("GFA: duplicate or zero segment id");
        if (s.seq.empty() || s.seq == "*") throw std::runtime_error("GFA: segment without sequence");
        seen[s.id] = 1;
        g.seq[s.id] = s.seq;
    }
    // all segments first, then all links in file order (HashGraph::from_gfa)
    for (auto& l : links) {
        if (l.from == 0 || l.from > max_id || l.to == 0 || l.to > max_id)
            throw std::runtime_error("GFA: link references a missing segment");
        g.create_edge(make_handle(l.from, l.frev), make_handle(l.to, l.trev));
    }
    return g;
}

HashGraph graph_from_gfa_file(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    if (!in) throw std::runtime_error("cannot open GFA file: " + path);
    std::stringstream ss;
    ss << in.rdbuf();
    return graph_from_gfa_text(ss.str());
}

// ---------------------------------------------------------------------------------------------
// k-mer enumeration (src/kmer.rs:347-505)
// ---------------------------------------------------------------------------------------------

namespace {

struct GraphKmer {
    std::string seq;
    bool begin_rev; uint64_t begin;   // begin_offset
    bool end_rev;   uint64_t end;     // end_offset
    handle_t first, last;
    bool handle_orient;
    uint64_t forks;
    bool operator==(const GraphKmer& o) const {
        return seq == o.seq && begin_rev == o.begin_rev && begin == o.begin && end_rev == o.end_rev &&
               end == o.end && first == o.first && last == o.last && handle_orient == o.handle_orient &&
               forks == o.forks;
    }
};

// first n bytes of g.sequence(h) without materialising a long reverse complement
std::string seq_prefix(const HashGraph& g, handle_t h, uint64_t n) {
    const std::string& s = g.seq[handle_id(h)];
    uint64_t m = std::min<uint64_t>(n, s.size());
    if (!handle_rev(h)) return s.substr(0, m);
    std::string out(m, 'N');
    for (uint64_t i = 0; i < m; i++) out[i] = hg_comp(s[s.size() - 1 - i]);
    return out;
}

void kmers_from_handle_orient(const HashGraph& g, handle_t handle_in, bool orient, uint64_t k,
                              uint64_t edge_max, uint64_t degree_max, std::vector<GraphKmer>& out_all) {
    handle_t handle = handle_in;
    std::vector<GraphKmer> complete;
    const std::vector<handle_t> nbrs0 = g.edges_right(handle);
    if (nbrs0.size() > degree_max) return;                                   // :361-372
    const std::string hseq = g.sequence(handle);
    const uint64_t hlen = hseq.size();
    std::vector<GraphKmer> stack;
    for (uint64_t i = 0; i < hlen; i++) {                                    // :383
        uint64_t end = std::min(i + k, hlen);
        GraphKmer km{hseq.substr(i, end - i), handle_rev(handle), i, handle_rev(handle), end,
                     handle, handle, orient, 0};
        if (km.seq.find('N') != std::string::npos) return;                   // :401-403 drops everything
        if (km.seq.size() == k) {
            complete.push_back(std::move(km));                               // :408-411
        } else {
            uint64_t next_count = nbrs0.size();
            if (next_count < degree_max || km.forks < edge_max) {            // :427-430
                for (handle_t nb : nbrs0) {
                    GraphKmer inc = km;
                    inc.last = nb;
                    if (next_count > 1) inc.forks += 1;
                    stack.push_back(std::move(inc));
                }
            }
        }
    }
    while (!stack.empty()) {                                                 // :449 LIFO
        GraphKmer c = std::move(stack.back());
        stack.pop_back();
        handle = c.last;
        const uint64_t hl = g.seq[handle_id(handle)].size();
        uint64_t end = std::min<uint64_t>(k - c.seq.size(), hl);
        std::string add = seq_prefix(g, handle, end);
        c.seq += add;                                                        // extend_kmer :80-84
        c.end_rev = handle_rev(handle);
        c.end = add.size();
        c.last = handle;
        if (c.seq.find('N') != std::string::npos) return;                    // :459-461
        if (c.seq.size() == k) {
            complete.push_back(std::move(c));
        } else {
            const std::vector<handle_t> nbrs = g.edges_right(handle);
            for (handle_t nb : nbrs) {                                       // :472-495
                uint64_t next_count = nbrs.size();
                if (next_count < degree_max || c.forks < edge_max) {
                    GraphKmer inc = c;
                    inc.last = nb;
                    if (next_count > 1) inc.forks += 1;
                    stack.push_back(std::move(inc));
                }
            }
        }
    }
    for (auto& x : complete) out_all.push_back(std::move(x));
}

}  // namespace

FlatIndex build_index(const HashGraph& g, uint32_t k, uint64_t max_furcations, uint64_t max_degree) {
    if (k == 0 || k > 32) throw std::runtime_error("kmer length must be in 1..32 for the B200 path");
    FlatIndex ix;
    ix.k = k;
    ix.n_nodes = g.n_nodes;
    // ---- find_forward_sequence (src/utils.rs:81-146): handles sorted by id
    for (uint64_t id = 1; id <= g.n_nodes; id++) {
        handle_t h = make_handle(id, false);
        std::vector<handle_t> l = g.edges_left(h), r = g.edges_right(h);
        ix.node_ref.push_back(ix.seq_fwd.size());
        ix.node_ref.push_back(ix.edges.size());
        ix.node_ref.push_back(l.size());
        ix.edges.insert(ix.edges.end(), l.begin(), l.end());
        ix.edges.insert(ix.edges.end(), r.begin(), r.end());
        ix.seq_fwd += g.seq[id];
    }
    ix.node_ref.push_back(ix.seq_fwd.size());
    ix.node_ref.push_back(ix.edges.size());
    ix.node_ref.push_back(0);
    ix.seq_len = ix.seq_fwd.size();
    ix.n_edges = ix.edges.size();
    ix.seq_rev = reverse_complement(ix.seq_fwd);

    // ---- generate_kmers_parallel (src/kmer.rs:277-304)
    std::vector<GraphKmer> kmers;
    for (uint64_t id = 1; id <= g.n_nodes; id++) {
        handle_t fh = make_handle(id, false);
        kmers_from_handle_orient(g, fh, true, k, max_furcations, max_degree, kmers);
        kmers_from_handle_orient(g, handle_flip(fh), false, k, max_furcations, max_degree, kmers);
    }
    std::stable_sort(kmers.begin(), kmers.end(),
                     [](const GraphKmer& a, const GraphKmer& b) { return a.seq < b.seq; });
    kmers.erase(std::unique(kmers.begin(), kmers.end()), kmers.end());   // consecutive dedup, all fields

    if (kmers.empty()) return ix;   // the reference would panic at src/kmer.rs:828 (`last().unwrap()`)

    // ---- generate_pos_on_ref_2 (src/kmer.rs:816-928)
    struct KP { uint64_t so, sp, eo, ep; };
    std::vector<std::vector<KP>> groups;
    std::vector<std::string> keys;
    std::vector<KP> cur;
    const GraphKmer final_kmer = kmers.back();
    bool have_last = false;
    std::string last_seq;
    const uint64_t L = ix.seq_len;
    auto seq_pos = [&](handle_t h) -> uint64_t {                           // get_seq_pos :752-770
        uint64_t start = ix.seq_idx(handle_id(h) - 1);
        uint64_t hl = g.seq[handle_id(h)].size();
        return handle_rev(h) ? L - start - hl : start;
    };
    for (const GraphKmer& km : kmers) {
        KP p{km.begin_rev ? 1ull : 0ull, seq_pos(km.first) + km.begin, km.end_rev ? 1ull : 0ull,
             seq_pos(km.last) + km.end};
        if (!have_last) {
            cur.push_back(p);
            last_seq = km.seq;
            have_last = true;
        } else if (last_seq != km.seq) {
            keys.push_back(last_seq);
            groups.push_back(cur);
            cur.assign(1, p);
            last_seq = km.seq;
        } else {
            cur.push_back(p);
        }
        if (km == final_kmer) {            // :885-889 (literal: fires on any element equal to the last)
            keys.push_back(km.seq);
            groups.push_back(cur);
        }
    }
    for (size_t i = 1; i < keys.size(); i++)
        if (keys[i] == keys[i - 1])
            throw std::runtime_error(
                "index build: the reference's final-kmer flush (src/kmer.rs:885) fired twice; it would "
                "hand duplicate keys to boomphf");
    for (auto& grp : groups)
        std::sort(grp.begin(), grp.end(), [](const KP& a, const KP& b) {     // derived Ord of KmerPos
            if (a.so != b.so) return a.so < b.so;
            if (a.sp != b.sp) return a.sp < b.sp;
            if (a.eo != b.eo) return a.eo < b.eo;
            return a.ep < b.ep;
        });
    for (size_t i = 0; i < groups.size(); i++) {
        uint64_t packed;
        if (!pack_kmer(keys[i].data(), k, &packed))
            throw std::runtime_error("index k-mer outside uppercase ACGT cannot be 2-bit packed: " + keys[i]);
        ix.key_kmer.push_back(keys[i]);
        ix.key_kmer2bit.push_back(packed);
        ix.key_offset.push_back(ix.tab_start.size());
        for (const KP& p : groups[i]) {
            ix.tab_start.push_back(p.sp | (p.so ? POS_REV_BIT : 0));
            ix.tab_end.push_back(p.ep | (p.eo ? POS_REV_BIT : 0));
        }
        ix.tab_start.push_back(POS_DELIM);
        ix.tab_end.push_back(POS_DELIM);
    }
    ix.n_kmer_pos = ix.tab_start.size();
    return ix;
}

// ---------------------------------------------------------------------------------------------
// .vgidx container
// ---------------------------------------------------------------------------------------------

namespace {
const char MAGIC[8] = {'V', 'G', 'I', 'D', 'X', '0', '0', '1'};
template <class T> void wr(std::ostream& o, const T& v) { o.write((const char*)&v, sizeof(T)); }
template <class T> void rd(std::istream& i, T& v) { i.read((char*)&v, sizeof(T)); }
void wr_vec(std::ostream& o, const std::vector<uint64_t>& v) {
    uint64_t n = v.size(); wr(o, n); o.write((const char*)v.data(), n * 8);
}
void rd_vec(std::istream& i, std::vector<uint64_t>& v) {
    uint64_t n; rd(i, n); v.resize(n); i.read((char*)v.data(), n * 8);
}
void wr_str(std::ostream& o, const std::string& s) { uint64_t n = s.size(); wr(o, n); o.write(s.data(), n); }
void rd_str(std::istream& i, std::string& s) { uint64_t n; rd(i, n); s.resize(n); i.read(&s[0], n); }
}  // namespace

void save_index(const FlatIndex& ix, const std::string& path) {
    std::ofstream o(path, std::ios::binary);
    if (!o) throw std::runtime_error("cannot write " + path);
    o.write(MAGIC, 8);
    uint64_t k = ix.k;
    wr(o, k); wr(o, ix.seq_len); wr(o, ix.n_nodes); wr(o, ix.n_edges); wr(o, ix.n_kmer_pos);
    wr_str(o, ix.seq_fwd);
    wr_vec(o, ix.node_ref); wr_vec(o, ix.edges); wr_vec(o, ix.key_kmer2bit); wr_vec(o, ix.key_offset);
    wr_vec(o, ix.tab_start); wr_vec(o, ix.tab_end);
}

static void unpack_kmer(uint64_t v, uint32_t k, std::string& s) {
    s.assign(k, 'A');
    for (uint32_t i = 0; i < k; i++) s[k - 1 - i] = "ACGT"[(v >> (2 * i)) & 3];
}

FlatIndex load_index(const std::string& path) {
    std::ifstream i(path, std::ios::binary);
    if (!i) throw std::runtime_error("cannot open " + path);
    char m[8]; i.read(m, 8);
    if (memcmp(m, MAGIC, 8) != 0) throw std::runtime_error("not a .vgidx file: " + path);
    FlatIndex ix;
    uint64_t k;
    rd(i, k); ix.k = (uint32_t)k; rd(i, ix.seq_len); rd(i, ix.n_nodes); rd(i, ix.n_edges); rd(i, ix.n_kmer_pos);
    rd_str(i, ix.seq_fwd);
    rd_vec(i, ix.node_ref); rd_vec(i, ix.edges); rd_vec(i, ix.key_kmer2bit); rd_vec(i, ix.key_offset);
    rd_vec(i, ix.tab_start); rd_vec(i, ix.tab_end);
    if (!i) throw std::runtime_error("truncated .vgidx file: " + path);
    ix.seq_rev = reverse_complement(ix.seq_fwd);
    ix.key_kmer.resize(ix.key_kmer2bit.size());
    for (size_t j = 0; j < ix.key_kmer.size(); j++) unpack_kmer(ix.key_kmer2bit[j], ix.k, ix.key_kmer[j]);
    return ix;
}

}  // namespace vga
