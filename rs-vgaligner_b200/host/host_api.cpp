// libvga_host.so — see vga_host.h.  Pure host code (no CUDA, no oracle/).
#include "vga_host.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "flat_index.h"

using namespace vga;

static thread_local std::string g_err;
static void set_err(const std::string& s) { g_err = s; }

struct vgah_index {
    FlatIndex ix;
    std::string keys;            // n_keys * k bytes
};

struct vgah_reads {
    std::vector<std::string> names;
    std::string seqs;
    std::vector<uint64_t> off;   // n+1
    std::string names_joined;
    void finish() {
        names_joined.clear();
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

vgah_index_t* vgah_index_load(const char* path) {
    try { return wrap(load_index(path)); } catch (std::exception& e) { set_err(e.what()); return nullptr; }
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

vgah_reads_t* vgah_reads_simulate(const vgah_index_t* h, uint32_t n, uint32_t len, double sub, double ins, double del,
                                  double rev_frac, uint64_t seed) {
    const FlatIndex& ix = h->ix;
    if (ix.seq_len == 0) { set_err("empty graph"); return nullptr; }
    // position -> node id
    std::vector<uint32_t> p2n(ix.seq_len);
    for (uint64_t id = 1; id <= ix.n_nodes; id++)
        for (uint64_t p = ix.seq_idx(id - 1); p < ix.seq_idx(id); p++) p2n[p] = (uint32_t)id;
    Rng rng(seed * 0x9E3779B97F4A7C15ull + 12345);
    vgah_reads* r = new vgah_reads();
    r->off.push_back(0);
    const char* ACGT = "ACGT";
    const uint64_t want_src = (uint64_t)len + len / 2 + 16;    // template long enough to survive deletions
    std::string tmpl, read;
    for (uint32_t i = 0; i < n; i++) {
        for (int attempt = 0;; attempt++) {
            if (attempt > 1000) { set_err("read simulation cannot find a long enough walk"); delete r; return nullptr; }
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
                std::vector<uint64_t> outs;
                for (uint64_t x = b; x < e; x++) if (!(ix.edges[x] & 1)) outs.push_back(ix.edges[x] >> 1);
                if (outs.empty()) { dead = true; break; }
                id = outs[rng.below(outs.size())];
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
        r->names.push_back("read" + std::to_string(i));
        r->seqs += read;
        r->off.push_back(r->seqs.size());
    }
    r->finish();
    return r;
}

void vgah_reads_free(vgah_reads_t* r) { delete r; }
uint32_t vgah_reads_count(const vgah_reads_t* r) { return (uint32_t)r->names.size(); }
const char* vgah_reads_seqs(const vgah_reads_t* r) { return r->seqs.data(); }
const uint64_t* vgah_reads_offsets(const vgah_reads_t* r) { return r->off.data(); }
const char* vgah_reads_names(const vgah_reads_t* r) { return r->names_joined.c_str(); }

int vgah_reads_write_fasta(const vgah_reads_t* r, const char* path) {
    std::ofstream o(path);
    if (!o) { set_err(std::string("cannot write ") + path); return -1; }
    for (size_t i = 0; i < r->names.size(); i++)
        o << '>' << r->names[i] << '\n' << r->seqs.substr(r->off[i], r->off[i + 1] - r->off[i]) << '\n';
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

// ---- GAF writer ---------------------------------------------------------------------------------
char* vgah_gaf(const vgah_index_t* h, const VgaResultsView* R, const vgah_reads_t* reads, int which, uint64_t* len) {
    const FlatIndex& ix = h->ix;
    std::string out;
    // rank / select over seq_bv (src/index.rs:427-480) from the node table
    std::vector<uint64_t> starts(ix.n_nodes);
    for (uint64_t i = 0; i < ix.n_nodes; i++) starts[i] = ix.seq_idx(i);
    auto node_of = [&](uint64_t pos) { return (uint64_t)(std::upper_bound(starts.begin(), starts.end(), pos) - starts.begin()); };
    auto label = [&](uint64_t hd) {
        uint64_t id = hd >> 1, b = ix.seq_idx(id - 1), e = ix.seq_idx(id);
        return (hd & 1) ? ix.seq_rev.substr(ix.seq_len - e, e - b) : ix.seq_fwd.substr(b, e - b);
    };
    for (uint32_t r = 0; r < R->n_reads; r++) {
        const std::string& name = reads->names[r];
        const uint64_t qlen = reads->off[r + 1] - reads->off[r];
        auto placeholder = [&]() { out += name + "\t" + std::to_string(qlen) + "\t*\t*\t*\t*\t*\t*\t*\t*\t*\t0\t*\n"; };
        if (which == 0) {
            if (R->chain_off[r] == R->chain_off[r + 1]) { placeholder(); continue; }
            for (uint64_t c = R->chain_off[r]; c < R->chain_off[r + 1]; c++) {
                const uint64_t b = R->chain_anchor_off[c], e = R->chain_anchor_off[c + 1];
                std::string path;
                for (uint64_t a = b; a < e; a++) {
                    const uint64_t tb = R->ca_tb[a], tei = (uint64_t)R->ca_te[a] - 1;
                    const uint64_t n1 = node_of(tb), n2 = node_of(tei);
                    path += "(>" + std::to_string(n1) + ":" + std::to_string(tb - starts[n1 - 1]) + ",>" + std::to_string(n2) +
                            ":" + std::to_string(tei - starts[n2 - 1]) + "),";
                }
                out += name + "\t" + std::to_string(qlen) + "\t" + std::to_string(R->ca_qb[b]) + "\t" +
                       std::to_string(R->ca_qb[e - 1] + ix.k) + "\t+\t" + path + "\t0\t0\t0\t0\t0\t0\t" +
                       "ta:Z:chain,n_anchors: " + std::to_string(e - b) + "\n";
            }
        } else {
            if (R->aln_status[r] == VGA_ALN_PLACEHOLDER) { placeholder(); continue; }
            std::string path, cs = "cs:Z:", cg;
            for (uint64_t p = R->aln_path_off[r]; p < R->aln_path_off[r + 1]; p++)
                path += ((R->aln_path[p] & 1) ? "<" : ">") + std::to_string(R->aln_path[p] >> 1);
            const char* q = reads->seqs.data() + reads->off[r];
            uint64_t pi = R->aln_path_off[r], qi = 0, off = R->aln_start_off[r];
            std::string node_seq = pi < R->aln_path_off[r + 1] ? label(R->aln_path[pi]) : std::string();
            auto next_base = [&]() -> char {
                while (off >= node_seq.size()) { pi++; node_seq = label(R->aln_path[pi]); off = 0; }
                return node_seq[off++];
            };
            uint32_t mrun = 0;
            auto flush_m = [&]() { if (mrun) { cg += std::to_string(mrun) + "M"; mrun = 0; } };
            for (uint64_t t = R->aln_run_off[r]; t < R->aln_run_off[r + 1]; t++) {
                const uint32_t op = R->aln_runs[t] >> 28, n = R->aln_runs[t] & 0x0fffffffu;
                if (op == 0) { cs += ":" + std::to_string(n); for (uint32_t x = 0; x < n; x++) { next_base(); qi++; } mrun += n; }
                else if (op == 1) { for (uint32_t x = 0; x < n; x++) { cs += "*"; cs += (char)tolower(next_base()); cs += (char)tolower(q[qi++]); } mrun += n; }
                else if (op == 2) { flush_m(); cs += "-"; for (uint32_t x = 0; x < n; x++) cs += (char)tolower(next_base()); cg += std::to_string(n) + "D"; }
                else { flush_m(); cs += "+"; for (uint32_t x = 0; x < n; x++) cs += (char)tolower(q[qi++]); cg += std::to_string(n) + "I"; }
            }
            flush_m();
            out += name + "\t" + std::to_string(qlen) + "\t0\t" + std::to_string(qlen) + "\t+\t" + path + "\t" +
                   std::to_string(R->aln_path_len[r]) + "\t" + std::to_string(R->aln_start_off[r]) + "\t" +
                   std::to_string(R->aln_end_off[r]) + "\t0\t" + std::to_string(R->aln_n_aligned[r]) + "\t255\t" + "as:i:-30 " +
                   cs + ",cg:Z:" + cg + "\n";
        }
    }
    char* c = (char*)malloc(out.size() + 1);
    memcpy(c, out.data(), out.size() + 1);
    if (len) *len = out.size();
    return c;
}

}  // extern "C"
