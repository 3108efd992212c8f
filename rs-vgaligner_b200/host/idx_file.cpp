// `.idx` reader (and a writer for round trips) — the file `vgaligner index` writes: bincode::serialize(&Index)
// (reference src/index.rs:284-305, src/serialization.rs:34-54).  bincode 1.3.3 defaults fix the layout of everything the
// reference declares itself: little endian, fixed-width integers, u64 lengths, enum tags u32, Option tags u8, bool u8, fields
// in declaration order (src/index.rs:37-90):
//   kmer_length u64 | seq_length u64 | seq_fwd (len + bytes) | seq_rev (len + bytes) | seq_bv [bv 0.11.1] | n_edges u64 |
//   edges (len + u64 each; custom serializer src/serialization.rs:17-30) | n_nodes u64 | node_ref (len + 3 x u64 each,
//   src/utils.rs:14-22) | n_kmers u64 | n_kmer_pos u64 | bhpf [boomphf 0.5.9] | kmer_pos_ref (len + u64 each) |
//   sampling_rate (u8 tag [+ u64]) | kmer_pos_table (len + {u32 orient, u64 position, u32 orient, u64 position} each,
//   src/kmer.rs:18-31, 732-738) | loaded u8
// The two bracketed fields are serialised by third-party crates whose sources are not in /root/reference; their inner
// layouts below are RECALLED and UNVERIFIED (no cargo in this image, no real `.idx` to read).  The reader therefore does
// not depend on them:
//   * seq_bv is not needed (NodeRef.seq_idx holds the node starts).  The reader tries the recalled layout and otherwise
//     re-synchronises on the fields behind it, which validate each other: n_edges followed by a vector of that length,
//     n_nodes followed by a vector of n_nodes + 1 NodeRefs whose first is (0, 0, .) and whose last is (seq_length, n_edges, 0).
//   * bhpf is not needed either: kmer_pos_ref is in the order of the table's groups (src/kmer.rs:894-923), so the offset of
//     key i is where the i-th delimiter-terminated run starts.  Everything behind bhpf is located from the END of the file
//     (its sizes are given by n_kmers and n_kmer_pos, which precede bhpf).
// The keys of a `.idx` are ahash values of the k-mer strings; the 2-bit keys of the device table are derived from the table
// itself (vgah_derive_keys with no hash function: key_derive.cpp).  Sampled indexes (sampling_rate = Some) are rejected like
// everywhere else on this path.
#include <cstring>
#include <fstream>
#include <stdexcept>

#include "flat_index.h"
#include "vga_host.h"

namespace vga {

namespace {

struct Cur {
    const uint8_t* p;
    size_t n, at;
    bool has(size_t b) const { return at + b <= n; }
    uint64_t u64_at(size_t o) const { uint64_t v; memcpy(&v, p + o, 8); return v; }
    uint64_t u64() { if (!has(8)) throw std::runtime_error(".idx: truncated"); uint64_t v = u64_at(at); at += 8; return v; }
    std::string str() {
        const uint64_t len = u64();
        if (!has(len)) throw std::runtime_error(".idx: truncated string");
        std::string s((const char*)p + at, len);
        at += len;
        return s;
    }
};

// [RECALLED, bv 0.11.1 with its serde feature] BitVec<usize> { bits: Inner(Option<Box<[usize]>>), len: u64 }
//   -> u8 Option tag, (u64 block count, u64 blocks; bit i = bit i % 64 of block i / 64), u64 len
bool skip_bitvec_recalled(Cur& c, uint64_t seq_len) {
    Cur t = c;
    if (!t.has(1)) return false;
    const uint8_t tag = t.p[t.at++];
    if (tag > 1) return false;
    if (tag == 1) {
        if (!t.has(8)) return false;
        const uint64_t nb = t.u64();
        if (nb > (seq_len + 1 + 63) / 64 + 1 || !t.has(nb * 8)) return false;
        t.at += nb * 8;
    }
    if (!t.has(8)) return false;
    const uint64_t len = t.u64();
    if (len != seq_len + 1) return false;
    c = t;
    return true;
}

// n_edges | edges | n_nodes | node_ref starting at `at`?  (the fields validate each other)
bool graph_block_at(const Cur& c, size_t at, uint64_t seq_len, FlatIndex* out, size_t* end) {
    Cur t = c;
    t.at = at;
    if (!t.has(16)) return false;
    const uint64_t n_edges = t.u64(), elen = t.u64();
    if (elen != n_edges || !t.has(elen * 8 + 16)) return false;
    const size_t edges_at = t.at;
    t.at += elen * 8;
    const uint64_t n_nodes = t.u64(), nlen = t.u64();
    if (nlen != n_nodes + 1 || n_nodes == 0 || !t.has(nlen * 24)) return false;
    const size_t nr_at = t.at;
    if (t.u64_at(nr_at) != 0 || t.u64_at(nr_at + 8) != 0) return false;
    const size_t last = nr_at + (nlen - 1) * 24;
    if (t.u64_at(last) != seq_len || t.u64_at(last + 8) != n_edges || t.u64_at(last + 16) != 0) return false;
    if (out) {
        out->n_edges = n_edges; out->n_nodes = n_nodes;
        out->edges.resize(elen);
        memcpy(out->edges.data(), t.p + edges_at, elen * 8);
        out->node_ref.resize(nlen * 3);
        memcpy(out->node_ref.data(), t.p + nr_at, nlen * 24);
    }
    *end = nr_at + nlen * 24;
    return true;
}

void put_u64(std::string& o, uint64_t v) { o.append((const char*)&v, 8); }
void put_u32(std::string& o, uint32_t v) { o.append((const char*)&v, 4); }

}  // namespace

FlatIndex load_idx(const std::string& path) {
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) throw std::runtime_error("cannot open " + path);
    const size_t size = (size_t)f.tellg();
    std::string buf(size, '\0');
    f.seekg(0);
    f.read(&buf[0], size);
    if (!f) throw std::runtime_error("cannot read " + path);
    Cur c{(const uint8_t*)buf.data(), size, 0};
    FlatIndex ix;
    const uint64_t k = c.u64();
    if (k == 0 || k > 32) throw std::runtime_error(".idx: kmer_length outside 1..32 (not an index of this format, or unsupported)");
    ix.k = (uint32_t)k;
    ix.seq_len = c.u64();
    ix.seq_fwd = c.str();
    ix.seq_rev = c.str();
    if (ix.seq_fwd.size() != ix.seq_len || ix.seq_rev.size() != ix.seq_len) throw std::runtime_error(".idx: linearisation length mismatch");
    // ---- seq_bv (skipped), then the graph block
    size_t graph_end = 0;
    bool ok = false;
    {
        Cur t = c;
        if (skip_bitvec_recalled(t, ix.seq_len)) ok = graph_block_at(c, t.at, ix.seq_len, &ix, &graph_end);
    }
    if (!ok) {
        // the recalled bit-vector layout does not hold: find the graph block by its self-validating fields
        const size_t limit = std::min(size, c.at + (ix.seq_len / 8 + 4096) * 2);
        for (size_t at = c.at; at + 16 <= limit && !ok; at++) ok = graph_block_at(c, at, ix.seq_len, &ix, &graph_end);
    }
    if (!ok) throw std::runtime_error(".idx: cannot locate n_edges / edges / n_nodes / node_ref behind seq_bv");
    c.at = graph_end;
    const uint64_t n_kmers = c.u64(), n_kmer_pos = c.u64();
    // ---- everything behind bhpf, from the end of the file
    if (size < 1 + 8 + n_kmer_pos * 24) throw std::runtime_error(".idx: truncated (kmer_pos_table)");
    const size_t tab_at = size - 1 - n_kmer_pos * 24;
    if (c.u64_at(tab_at - 8) != n_kmer_pos) throw std::runtime_error(".idx: kmer_pos_table length field does not match n_kmer_pos");
    size_t opt_at = tab_at - 8 - 1;                                   // sampling_rate = None: one tag byte
    bool sampled = false;
    auto ref_ok = [&](size_t opt) {
        if (opt < 8 + n_kmers * 8 + c.at) return false;
        const size_t ref_at = opt - n_kmers * 8;
        return c.u64_at(ref_at - 8) == n_kmers;
    };
    if (!(c.p[opt_at] == 0 && ref_ok(opt_at))) {
        opt_at = tab_at - 8 - 9;                                      // Some(rate): tag + u64
        if (opt_at < size && c.p[opt_at] == 1 && ref_ok(opt_at)) sampled = true;
        else throw std::runtime_error(".idx: cannot locate kmer_pos_ref / sampling_rate in front of kmer_pos_table");
    }
    if (sampled) throw std::runtime_error(".idx: sampled indexes (-r) are not supported on the GPU path");
    // ---- table: orientation in bit 63, one delimiter per key (src/kmer.rs:740-749)
    ix.tab_start.resize(n_kmer_pos); ix.tab_end.resize(n_kmer_pos);
    ix.key_offset.reserve(n_kmers);
    bool at_start = true;
    for (uint64_t i = 0; i < n_kmer_pos; i++) {
        const uint8_t* e = c.p + tab_at + i * 24;
        uint32_t so, eo; uint64_t sp, ep;
        memcpy(&so, e, 4); memcpy(&sp, e + 4, 8); memcpy(&eo, e + 12, 4); memcpy(&ep, e + 16, 8);
        if (so > 1 || eo > 1) throw std::runtime_error(".idx: bad SeqOrient tag in kmer_pos_table");
        const bool delim = so == 1 && eo == 1 && sp == ~0ull && ep == ~0ull;
        if (delim) { ix.tab_start[i] = POS_DELIM; ix.tab_end[i] = POS_DELIM; at_start = true; continue; }
        if (at_start) { ix.key_offset.push_back(i); at_start = false; }
        ix.tab_start[i] = sp | (so ? POS_REV_BIT : 0);
        ix.tab_end[i] = ep | (eo ? POS_REV_BIT : 0);
    }
    if (ix.key_offset.size() != n_kmers) throw std::runtime_error(".idx: the number of delimiter-terminated runs differs from n_kmers");
    ix.n_kmer_pos = n_kmer_pos;
    // ---- the 2-bit keys, from the table (no ahash outside Rust)
    ix.key_kmer2bit.resize(n_kmers);
    VgaIndexView v;
    memset(&v, 0, sizeof(v));
    v.k = ix.k; v.seq_len = ix.seq_len; v.n_nodes = ix.n_nodes; v.n_edges = ix.n_edges; v.n_keys = n_kmers; v.n_table = n_kmer_pos;
    v.seq_fwd = (const uint8_t*)ix.seq_fwd.data(); v.node_ref = ix.node_ref.data(); v.edges = ix.edges.data();
    v.key_offset = ix.key_offset.data(); v.tab_start = ix.tab_start.data(); v.tab_end = ix.tab_end.data();
    const int rc = vgah_derive_keys(&v, nullptr, nullptr, nullptr, ix.key_kmer2bit.data(), nullptr, nullptr);
    if (rc != 0) throw std::runtime_error(".idx: the k-mer keys cannot be derived from the table (vgah_derive_keys " + std::to_string(rc) + ")");
    ix.key_kmer.resize(n_kmers);
    for (size_t j = 0; j < n_kmers; j++) {
        ix.key_kmer[j].assign(ix.k, 'A');
        for (uint32_t t = 0; t < ix.k; t++) ix.key_kmer[j][ix.k - 1 - t] = "ACGT"[(ix.key_kmer2bit[j] >> (2 * t)) & 3];
    }
    return ix;
}

// Writer for round trips and for handing an index to a machine that has the Rust binary.  key_hash: the keys to store in
// kmer_pos_ref (a real file holds ahash(k-mer); any 64-bit values serve the reader, which never interprets them).  The bhpf
// field is written as an EMPTY minimal perfect hash in the recalled boomphf layout (Mphf { bitvecs: [], ranks: [] }) followed by
// the value vector: a file from this writer is for this reader, the Rust side rebuilds its Mphf from kmer_pos_ref.
void save_idx(const FlatIndex& ix, const std::string& path, const std::vector<uint64_t>& key_hash) {
    if (key_hash.size() != ix.key_offset.size()) throw std::runtime_error("save_idx: one hash per key expected");
    std::string o;
    put_u64(o, ix.k); put_u64(o, ix.seq_len);
    put_u64(o, ix.seq_fwd.size()); o += ix.seq_fwd;
    put_u64(o, ix.seq_rev.size()); o += ix.seq_rev;
    {   // seq_bv [RECALLED layout]: a bit at every node start and at seq_len (src/utils.rs:130-135)
        const uint64_t nbits = ix.seq_len + 1, nb = (nbits + 63) / 64;
        std::vector<uint64_t> blocks(nb, 0);
        for (uint64_t i = 0; i <= ix.n_nodes; i++) { const uint64_t b = ix.seq_idx(i); blocks[b / 64] |= 1ull << (b % 64); }
        o.push_back(1);
        put_u64(o, nb);
        o.append((const char*)blocks.data(), nb * 8);
        put_u64(o, nbits);
    }
    put_u64(o, ix.n_edges);
    put_u64(o, ix.edges.size()); o.append((const char*)ix.edges.data(), ix.edges.size() * 8);
    put_u64(o, ix.n_nodes);
    put_u64(o, ix.n_nodes + 1); o.append((const char*)ix.node_ref.data(), ix.node_ref.size() * 8);
    put_u64(o, ix.key_offset.size()); put_u64(o, ix.tab_start.size());
    put_u64(o, 0); put_u64(o, 0);                                    // bhpf.mphf [RECALLED]: no bit vectors, no ranks
    put_u64(o, ix.key_offset.size()); o.append((const char*)ix.key_offset.data(), ix.key_offset.size() * 8);   // bhpf.values
    put_u64(o, key_hash.size()); o.append((const char*)key_hash.data(), key_hash.size() * 8);
    o.push_back(0);                                                  // sampling_rate: None
    put_u64(o, ix.tab_start.size());
    for (size_t i = 0; i < ix.tab_start.size(); i++) {
        const uint64_t s = ix.tab_start[i], e = ix.tab_end[i];
        if (s == POS_DELIM) { put_u32(o, 1); put_u64(o, ~0ull); put_u32(o, 1); put_u64(o, ~0ull); continue; }
        put_u32(o, (s & POS_REV_BIT) ? 1 : 0); put_u64(o, s & ~POS_REV_BIT);
        put_u32(o, (e & POS_REV_BIT) ? 1 : 0); put_u64(o, e & ~POS_REV_BIT);
    }
    o.push_back(0);                                                  // loaded: false
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot write " + path);
    f.write(o.data(), o.size());
}

}  // namespace vga
