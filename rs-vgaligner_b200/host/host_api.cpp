// libvga_host.so — see vga_host.h.  Pure host code (no CUDA, no oracle/).
#include "vga_host.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <fstream>
#include <mutex>
#include <thread>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "flat_index.h"

using namespace vga;

namespace vga {
struct GafTables;
const GafTables* gaf_tables_new(const FlatIndex& ix);
void gaf_tables_free(const GafTables* t);
char* gaf_format(const FlatIndex& ix, const GafTables& tb, const VgaResultsView* R, const char* names_joined, uint64_t name_base,
                 const char* seqs, const uint64_t* off, int which, int threads, uint64_t* len, uint64_t* n_failed);
}

static thread_local std::string g_err;
static void set_err(const std::string& s) { g_err = s; }

struct vgah_index {
    FlatIndex ix;
    std::string keys;            // n_keys * k bytes
    mutable std::once_flag gaf_once;
    mutable const GafTables* gaf_tables = nullptr;     // position -> node tables of the GAF writer, built on first use
    ~vgah_index() { if (gaf_tables) gaf_tables_free(gaf_tables); }
};

struct vgah_reads {
    bool auto_names = false;     // names are "read<i>" and are not materialised (simulated sets of millions of reads)
    std::vector<std::string> names;
    std::string seqs;
    std::vector<uint64_t> off;   // n+1
    std::string names_joined;
    void finish() {
        names_joined.clear();
        if (auto_names) return;
        for (size_t i = 0; i < names.size(); i++) { if (i) names_joined += '\n'; names_joined += names[i]; }
    }
};

static vgah_index* wrap(FlatIndex&& f) {
    vgah_index* h = new vgah_index();
    h->ix = std::move(f);
    h->keys.reserve(h->ix.key_kmer.size() * h->ix.k);
    for (auto& s : h->ix.key_kmer) h->keys += s;
    return h;
}

// ---- small deterministic RNG (splitmix64) ---------------------------------------------------
struct Rng {
    uint64_t s;
    explicit Rng(uint64_t seed) : s(seed) {}
    uint64_t next() {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    uint64_t below(uint64_t n) { return next() % n; }
    double uni() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
};

extern "C" {

const char* vgah_last_error(void) { return g_err.c_str(); }

vgah_index_t* vgah_index_build_file(const char* gfa_path, uint32_t k, uint64_t mf, uint64_t md) {
    try {
        HashGraph g = graph_from_gfa_file(gfa_path);
        return wrap(build_index(g, k, mf, md));
    } catch (std::exception& e) { set_err(e.what()); return nullptr; }
}

vgah_index_t* vgah_index_build_text(const char* gfa_text, uint32_t k, uint64_t mf, uint64_t md) {
    try {
        HashGraph g = graph_from_gfa_text(gfa_text);
        return wrap(build_index(g, k, mf, md));
    } catch (std::exception& e) { set_err(e.what()); return nullptr; }
}

static bool ends_with(const std::string& s, const char* suf) {
    const size_t n = strlen(suf);
    return s.size() >= n && s.compare(s.size() - n, n, suf) == 0;
}

// `.idx` = the reference's bincode file (Index::load_from_file, src/index.rs:296-305); anything else = this repository's container
vgah_index_t* vgah_index_load(const char* path) {
    try {
        const std::string p(path);
        return wrap(ends_with(p, ".idx") ? load_idx(p) : load_index(p));
    } catch (std::exception& e) { set_err(e.what()); return nullptr; }
}

int vgah_index_save_idx(const vgah_index_t* h, const char* path) {
    try {
        // stand-in for the ahash keys of a real file: FNV-1a of the k-mer strings (opaque to every reader)
        std::vector<uint64_t> hs(h->ix.key_kmer.size());
        for (size_t i = 0; i < hs.size(); i++) {
            uint64_t x = 0xcbf29ce484222325ull;
            for (char c : h->ix.key_kmer[i]) x = (x ^ (uint8_t)c) * 0x100000001b3ull;
            hs[i] = x;
        }
        save_idx(h->ix, path, hs);
        return 0;
    } catch (std::exception& e) { set_err(e.what()); return -1; }
}

int vgah_index_save(const vgah_index_t* h, const char* path) {
    try { save_index(h->ix, path); return 0; } catch (std::exception& e) { set_err(e.what()); return -1; }
}

void vgah_index_free(vgah_index_t* h) { delete h; }

int vgah_index_view(const vgah_index_t* h, VgaIndexView* v) {
    if (!h || !v) return -1;
    const FlatIndex& ix = h->ix;
    v->k = ix.k; v->seq_len = ix.seq_len; v->n_nodes = ix.n_nodes; v->n_edges = ix.n_edges;
    v->n_keys = ix.key_kmer2bit.size(); v->n_table = ix.tab_start.size();
    v->seq_fwd = (const uint8_t*)ix.seq_fwd.data(); v->node_ref = ix.node_ref.data(); v->edges = ix.edges.data();
    v->key_kmer2bit = ix.key_kmer2bit.data(); v->key_offset = ix.key_offset.data();
    v->tab_start = ix.tab_start.data(); v->tab_end = ix.tab_end.data();
    return 0;
}

const char* vgah_index_keys(const vgah_index_t* h, uint64_t* n_keys) {
    if (n_keys) *n_keys = h->ix.key_kmer.size();
    return h->keys.data();
}

const char* vgah_index_seq_rev(const vgah_index_t* h) { return h->ix.seq_rev.c_str(); }

// ---- reads ------------------------------------------------------------------------------------
vgah_reads_t* vgah_reads_from_file(const char* path) {
    try {
        std::string p(path);
        size_t dot = p.rfind('.');
        std::string ext = dot == std::string::npos ? "" : p.substr(dot + 1);
        bool fasta = ext == "fasta" || ext == "fa", fastq = ext == "fastq" || ext == "fq";
        if (!fasta && !fastq) throw std::runtime_error("Unrecognized file type");      // src/io.rs:86
        std::ifstream in(path);
        if (!in) throw std::runtime_error(std::string("cannot open ") + path);
        vgah_reads* r = new vgah_reads();
        r->off.push_back(0);
        std::string line;
        auto getl = [&](std::string& s) -> bool {
            if (!std::getline(in, s)) return false;
            if (!s.empty() && s.back() == '\r') s.pop_back();
            return true;
        };
        if (fasta) {        // one read per non-empty sequence line; repeated-name suffixing (src/io.rs:96-121)
            std::string last_name;
            int same = 0;
            while (getl(line)) {
                if (!line.empty() && line[0] == '>') { last_name = line.substr(1); same = 0; }
                else if (!line.empty()) {
                    r->names.push_back(same == 0 ? last_name : last_name + std::to_string(same));
                    r->seqs += line;
                    r->off.push_back(r->seqs.size());
                    same++;
                }
            }
        } else {            // strict 4-line records (src/io.rs:122-130)
            std::string n, s, a, b;
            while (getl(n) && getl(s) && getl(a) && getl(b)) {
                r->names.push_back(n.empty() ? n : n.substr(1));
                r->seqs += s;
                r->off.push_back(r->seqs.size());
            }
        }
        r->finish();
        return r;
    } catch (std::exception& e) { set_err(e.what()); return nullptr; }
}

// one simulated read: forward random walk from a uniform start base, error model, optional reverse complement.
// Returns false when no long enough walk was found in 1000 attempts.
static bool simulate_one(const FlatIndex& ix, const std::vector<uint32_t>& p2n, Rng& rng, uint32_t len, double sub, double ins,
                         double del, double rev_frac, std::string& tmpl, std::string& read) {
    const char* ACGT = "ACGT";
    const uint64_t want_src = (uint64_t)len + len / 2 + 16;    // template long enough to survive deletions
    for (int attempt = 0;; attempt++) {
        if (attempt > 1000) return false;
        tmpl.clear();
        uint64_t pos = rng.below(ix.seq_len);
        uint64_t id = p2n[pos];
        bool dead = false;
        while (tmpl.size() < want_src) {
            uint64_t end = ix.seq_idx(id);
            uint64_t take = std::min<uint64_t>(end - pos, want_src - tmpl.size());
            tmpl.append(ix.seq_fwd, pos, take);
            if (tmpl.size() >= want_src) break;
            // forward out-edges of the forward handle that lead to forward handles
            uint64_t b = ix.edge_idx(id - 1) + ix.edges_in(id - 1), e = ix.edge_idx(id);
            uint64_t n_out = 0;
            for (uint64_t x = b; x < e; x++) n_out += !(ix.edges[x] & 1);
            if (n_out == 0) { dead = true; break; }
            uint64_t pick = rng.below(n_out);
            for (uint64_t x = b; x < e; x++)
                if (!(ix.edges[x] & 1) && pick-- == 0) { id = ix.edges[x] >> 1; break; }
            pos = ix.seq_idx(id - 1);
        }
        if (dead && tmpl.size() < (uint64_t)len) continue;
        if (tmpl.find('N') != std::string::npos) continue;
        // error model
        read.clear();
        size_t t = 0;
        while (read.size() < len && t < tmpl.size()) {
            double u = rng.uni();
            if (u < ins) { read.push_back(ACGT[rng.below(4)]); continue; }
            if (u < ins + del) { t++; continue; }
            char c = tmpl[t++];
            if (u < ins + del + sub) { char d; do { d = ACGT[rng.below(4)]; } while (d == c); c = d; }
            read.push_back(c);
        }
        if (read.size() < len) continue;
        break;
    }
    if (rng.uni() < rev_frac) {
        std::string rc(read.size(), 'N');
        for (size_t x = 0; x < read.size(); x++) {
            char c = read[read.size() - 1 - x];
            rc[x] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : 'N';
        }
        read.swap(rc);
    }
    return true;
}

static void pos_to_node(const FlatIndex& ix, std::vector<uint32_t>& p2n) {
    p2n.resize(ix.seq_len);
    for (uint64_t id = 1; id <= ix.n_nodes; id++)
        for (uint64_t p = ix.seq_idx(id - 1); p < ix.seq_idx(id); p++) p2n[p] = (uint32_t)id;
}

vgah_reads_t* vgah_reads_simulate(const vgah_index_t* h, uint32_t n, uint32_t len, double sub, double ins, double del,
                                  double rev_frac, uint64_t seed) {
    const FlatIndex& ix = h->ix;
    if (ix.seq_len == 0) { set_err("empty graph"); return nullptr; }
    std::vector<uint32_t> p2n;
    pos_to_node(ix, p2n);
    Rng rng(seed * 0x9E3779B97F4A7C15ull + 12345);       // ONE stream for the whole set (read i depends on reads < i)
    vgah_reads* r = new vgah_reads();
    r->off.push_back(0);
    std::string tmpl, read;
    for (uint32_t i = 0; i < n; i++) {
        if (!simulate_one(ix, p2n, rng, len, sub, ins, del, rev_frac, tmpl, read)) {
            set_err("read simulation cannot find a long enough walk"); delete r; return nullptr;
        }
        r->names.push_back("read" + std::to_string(i));
        r->seqs += read;
        r->off.push_back(r->seqs.size());
    }
    r->finish();
    return r;
}

// Same generator with one RNG stream PER READ (seeded by (seed, first_id + i)), so that any sub-range of a read set can be
// produced by any thread or rank: used for the multi-million-read workloads (10 M x 150 bp, 100 k x 10 kb).  Every read has
// exactly `len` bases; names are "read<i>" and are not materialised.
vgah_reads_t* vgah_reads_simulate_mt(const vgah_index_t* h, uint64_t first_id, uint32_t n, uint32_t len, double sub, double ins,
                                     double del, double rev_frac, uint64_t seed, int threads) {
    const FlatIndex& ix = h->ix;
    if (ix.seq_len == 0) { set_err("empty graph"); return nullptr; }
    std::vector<uint32_t> p2n;
    pos_to_node(ix, p2n);
    vgah_reads* r = new vgah_reads();
    r->auto_names = true;
    r->seqs.assign((size_t)n * len, 'N');
    r->off.resize((size_t)n + 1);
    for (uint64_t i = 0; i <= n; i++) r->off[i] = i * len;
    int T = threads > 0 ? threads : (int)std::max(1u, std::thread::hardware_concurrency());
    if (n < 4096) T = 1;
    std::atomic<bool> failed{false};
    auto work = [&](int t) {
        std::string tmpl, read;
        const uint32_t a = (uint32_t)((uint64_t)n * t / T), b = (uint32_t)((uint64_t)n * (t + 1) / T);
        for (uint32_t i = a; i < b && !failed; i++) {
            Rng rng((seed * 0x9E3779B97F4A7C15ull + 12345) ^ ((first_id + i + 1) * 0xD6E8FEB86659FD93ull));
            if (!simulate_one(ix, p2n, rng, len, sub, ins, del, rev_frac, tmpl, read)) { failed = true; return; }
            memcpy(&r->seqs[(size_t)i * len], read.data(), len);
        }
    };
    if (T == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++) th.emplace_back(work, t);
        for (auto& x : th) x.join();
    }
    if (failed) { set_err("read simulation cannot find a long enough walk"); delete r; return nullptr; }
    return r;
}

void vgah_reads_free(vgah_reads_t* r) { delete r; }
uint32_t vgah_reads_count(const vgah_reads_t* r) { return (uint32_t)(r->off.size() - 1); }
const char* vgah_reads_seqs(const vgah_reads_t* r) { return r->seqs.data(); }
const uint64_t* vgah_reads_offsets(const vgah_reads_t* r) { return r->off.data(); }
const char* vgah_reads_names(const vgah_reads_t* r) {
    if (r->auto_names && r->names_joined.empty() && r->off.size() > 1) {      // materialise on demand
        vgah_reads* w = const_cast<vgah_reads*>(r);
        for (size_t i = 0; i + 1 < r->off.size(); i++) { if (i) w->names_joined += '\n'; w->names_joined += "read" + std::to_string(i); }
    }
    return r->names_joined.c_str();
}

int vgah_reads_write_fasta(const vgah_reads_t* r, const char* path) {
    std::ofstream o(path);
    if (!o) { set_err(std::string("cannot write ") + path); return -1; }
    for (size_t i = 0; i + 1 < r->off.size(); i++)
        o << '>' << (r->auto_names ? "read" + std::to_string(i) : r->names[i]) << '\n' << r->seqs.substr(r->off[i], r->off[i + 1] - r->off[i]) << '\n';
    return 0;
}

// ---- synthetic variation graph ------------------------------------------------------------------
char* vgah_synth_gfa(uint64_t backbone_len, uint32_t site_every, uint64_t seed, uint64_t* len) {
    Rng rng(seed * 0xD1B54A32D192ED03ull + 777);
    const char* ACGT = "ACGT";
    std::string out = "H\tVN:Z:1.0\n";
    std::vector<std::pair<uint64_t, uint64_t>> links;
    uint64_t next_id = 1, made = 0;
    auto rand_seq = [&](uint64_t n) { std::string s(n, 'A'); for (auto& c : s) c = ACGT[rng.below(4)]; return s; };
    auto add_node = [&](const std::string& s) { out += "S\t" + std::to_string(next_id) + "\t" + s + "\n"; return next_id++; };
    std::vector<uint64_t> tails;        // nodes whose out-edges go to the next backbone chunk
    while (made < backbone_len) {
        // backbone chunk: geometric-ish gap with mean site_every, at least 2 bp
        uint64_t gap = 2 + (uint64_t)(-std::log(1.0 - rng.uni()) * (site_every > 2 ? site_every - 2 : 1));
        gap = std::min<uint64_t>(gap, backbone_len - made);
        uint64_t chunk = add_node(rand_seq(gap));
        made += gap;
        for (uint64_t t : tails) links.push_back({t, chunk});
        tails.assign(1, chunk);
        if (made >= backbone_len) break;
        double u = rng.uni();
        if (u < 0.8) {                    // SNP bubble: two 1-bp alleles
            char a = ACGT[rng.below(4)], b;
            do { b = ACGT[rng.below(4)]; } while (b == a);
            uint64_t n1 = add_node(std::string(1, a)), n2 = add_node(std::string(1, b));
            links.push_back({chunk, n1}); links.push_back({chunk, n2});
            tails = {n1, n2};
            made += 1;
        } else {                          // insertion / deletion: optional 1-10 bp node with a bypass edge
            uint64_t l = 1 + rng.below(10);
            uint64_t nd = add_node(rand_seq(l));
            links.push_back({chunk, nd});
            tails = {chunk, nd};
            if (u >= 0.9) made += l;      // deletion: the node is part of the backbone
        }
    }
    std::sort(links.begin(), links.end());
    for (auto& l : links) out += "L\t" + std::to_string(l.first) + "\t+\t" + std::to_string(l.second) + "\t+\t0M\n";
    char* c = (char*)malloc(out.size() + 1);
    memcpy(c, out.data(), out.size() + 1);
    if (len) *len = out.size();
    return c;
}

void vgah_free_str(char* s) { free(s); }

// ---- GAF writer (gaf_writer.cpp) ---------------------------------------------------------------------
static const GafTables* tables_of(const vgah_index_t* h) {
    std::call_once(h->gaf_once, [h]() { h->gaf_tables = gaf_tables_new(h->ix); });
    return h->gaf_tables;
}

char* vgah_gaf_raw(const vgah_index_t* h, const VgaResultsView* R, const char* names_joined, uint64_t name_base, const char* seqs,
                   const uint64_t* seq_off, int which, int threads, uint64_t* len, uint64_t* n_failed) {
    if (!h || !R || !seq_off || (which < 0 || which > 3 || which == 2)) { set_err("vgah_gaf: bad argument"); return nullptr; }
    try {
        char* c = gaf_format(h->ix, *tables_of(h), R, names_joined, name_base, seqs, seq_off, which, threads, len, n_failed);
        if (!c) set_err("vgah_gaf: out of memory");
        return c;
    } catch (std::exception& e) { set_err(e.what()); return nullptr; }
}

char* vgah_gaf(const vgah_index_t* h, const VgaResultsView* R, const vgah_reads_t* reads, int which, uint64_t* len) {
    if (!reads) { set_err("vgah_gaf: bad argument"); return nullptr; }
    return vgah_gaf_raw(h, R, reads->auto_names ? nullptr : reads->names_joined.c_str(), 0, reads->seqs.data(), reads->off.data(), which, 0,
                        len, nullptr);
}

}  // extern "C"
