// Key derivation for the device k-mer table: what `Index::to_gpu` runs on the Rust side.
//
// A `.idx` keys its k-mer table by the ahash of the k-mer STRING (reference src/kmer.rs:931-934, src/index.rs:314-321);
// the strings themselves are not stored.  The device table is keyed by the 2-bit packed k-mer, so the host has to get
// the string of every key back from what the index does hold: the key's first table position (start, end), the
// linearisation, NodeRef and the edge lists.  `walk_kmer` re-walks the graph from `start` for k bases exactly like the
// k-mer enumeration did at build time (reference src/kmer.rs:347-505: bases of the first handle from the start offset,
// then the handles of handle_edges_iter(h, Right) in stored order) and keeps the walks that stop at `end`.
// (start, end) alone does not identify the k-mer — the two alleles of a SNP bubble give two k-mers with the same first
// and last position — so the candidates are told apart by the key's own hash: the caller passes its hash function
// (the Rust host: `generate_hash`; the tests: any 64-bit string hash, applied to the same strings when the keys were made).
// A host without the hash function (the C++ `.idx` reader: ahash is not reproducible outside Rust) passes hash = NULL and
// gets the same keys from the ORDER of the table instead (the rank rule at the end of vgah_derive_keys).
#include <algorithm>
#include <cstring>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "flat_index.h"
#include "vga_host.h"

namespace {

struct Walker {
    const VgaIndexView& v;
    std::string seq_rev;                 // reverse complement of seq_fwd (Index.seq_rev, reference src/index.rs:47)
    uint64_t L, N;
    explicit Walker(const VgaIndexView& view) : v(view), L(view.seq_len), N(view.n_nodes) {
        seq_rev.resize(L);
        for (uint64_t i = 0; i < L; i++) {
            const char c = (char)v.seq_fwd[L - 1 - i];
            char r;
            switch (c) {                 // reference src/dna.rs:5-33: acgtu / ACGTU complemented, everything else -> 'N'
                case 'A': r = 'T'; break; case 'C': r = 'G'; break; case 'G': r = 'C'; break; case 'T': case 'U': r = 'A'; break;
                case 'a': r = 't'; break; case 'c': r = 'g'; break; case 'g': r = 'c'; break; case 't': case 'u': r = 'a'; break;
                default: r = 'N';
            }
            seq_rev[i] = r;
        }
    }
    uint64_t seq_idx(uint64_t i) const { return v.node_ref[3 * i]; }
    uint64_t edge_idx(uint64_t i) const { return v.node_ref[3 * i + 1]; }
    uint64_t edges_in(uint64_t i) const { return v.node_ref[3 * i + 2]; }
    uint64_t node_len(uint64_t id) const { return seq_idx(id) - seq_idx(id - 1); }
    // node id (1-based) of a forward position: rank over the node starts (reference src/index.rs:427-439)
    uint64_t node_of_fwd(uint64_t pos) const {
        uint64_t lo = 0, hi = N;         // largest i with seq_idx(i) <= pos
        while (hi - lo > 1) { const uint64_t mid = (lo + hi) / 2; if (seq_idx(mid) <= pos) lo = mid; else hi = mid; }
        return lo + 1;
    }
    // first position of a handle on its own strand (reference src/kmer.rs:752-770)
    uint64_t seqpos(uint64_t id, bool rev) const { return rev ? L - seq_idx(id - 1) - node_len(id) : seq_idx(id - 1); }
    const char* bases(uint64_t id, bool rev) const { return rev ? seq_rev.data() + seqpos(id, true) : (const char*)v.seq_fwd + seq_idx(id - 1); }
    // handle_edges_iter(h, Right) in stored order (reference src/index.rs:584-606): forward handle = the outgoing part of
    // the node's run; reverse handle = the incoming part of the forward handle, flipped and reversed
    void outgoing(uint64_t id, bool rev, std::vector<uint64_t>& out) const {
        out.clear();
        const uint64_t b = edge_idx(id - 1), e = edge_idx(id), nin = edges_in(id - 1);
        if (!rev) for (uint64_t x = b + nin; x < e; x++) out.push_back(v.edges[x]);
        else for (uint64_t x = b + nin; x-- > b;) out.push_back(v.edges[x] ^ 1ull);
    }
    // all k-mers spelled by walks of k bases from (handle, offset) that stop at `end`.  Two walks that arrive at the same
    // handle with the same bases so far have the same future, so states (handle, prefix) are visited once: on the
    // microsatellite loops of 5-B3106 (1-bp nodes, out-degree 17, duplicated links) the plain depth-first walk the index
    // build does visits 17^10 paths for a few thousand distinct k-mers.
    bool walk(uint64_t id0, bool rev0, uint64_t off0, uint32_t k, uint64_t end_word, std::set<std::string>& found, size_t budget) const {
        std::set<std::pair<uint64_t, std::string>> seen;
        std::vector<std::pair<uint64_t, std::string>> stack;
        std::vector<uint64_t> nb;
        // the first handle is entered at off0; every later one at 0
        auto visit = [&](uint64_t h, uint64_t off, const std::string& prefix) {
            const uint64_t id = h >> 1;
            const bool rev = h & 1;
            const uint64_t len = node_len(id);
            const uint64_t take = std::min<uint64_t>(k - prefix.size(), len - off);
            std::string cur = prefix;
            cur.append(bases(id, rev) + off, take);
            if (cur.size() == k) {
                const uint64_t e = (seqpos(id, rev) + off + take) | (rev ? vga::POS_REV_BIT : 0);
                if (e == end_word) found.insert(cur);
                return;
            }
            outgoing(id, rev, nb);
            for (uint64_t t : nb)
                if (seen.insert({t, cur}).second) stack.push_back({t, cur});
        };
        visit(id0 * 2 + (rev0 ? 1 : 0), off0, std::string());
        while (!stack.empty()) {
            if (seen.size() > budget) return false;
            const std::pair<uint64_t, std::string> st = std::move(stack.back());
            stack.pop_back();
            visit(st.first, 0, st.second);
        }
        return true;
    }
};

}  // namespace

extern "C" int vgah_derive_keys(const VgaIndexView* view, const uint64_t* key_hash, vgah_hash_fn hash, void* user,
                                uint64_t* key_kmer2bit_out, uint64_t* n_ambiguous, uint64_t* n_hopping) {
    if (!view || (hash && !key_hash) || !key_kmer2bit_out || !view->seq_fwd || !view->node_ref || !view->key_offset ||
        !view->tab_start || !view->tab_end)
        return VGA_ERR_INVALID;
    const Walker W(*view);
    const uint32_t k = view->k;
    uint64_t amb = 0, hop = 0;
    std::set<std::string> found;
    std::vector<uint64_t> pending;                                            // keys left to the rank rule (no hash function given)
    std::map<std::pair<uint64_t, uint64_t>, std::vector<std::string>> cand;    // their candidate k-mers by first position pair
    for (uint64_t i = 0; i < view->n_keys; i++) {
        const uint64_t off = view->key_offset[i];
        if (off >= view->n_table) return VGA_ERR_INVALID;
        const uint64_t s = view->tab_start[off], e = view->tab_end[off];
        if (s == vga::POS_DELIM) return VGA_ERR_INVALID;            // a key without a position
        const bool rev = (s & vga::POS_REV_BIT) != 0;
        const uint64_t p = s & ~vga::POS_REV_BIT;
        if (p >= W.L) return VGA_ERR_INVALID;
        // the handle that holds the start position, and the offset inside it (reverse strand: mirrored coordinates)
        const uint64_t id = rev ? W.node_of_fwd(W.L - 1 - p) : W.node_of_fwd(p);
        const uint64_t o = p - W.seqpos(id, rev);
        found.clear();
        if (!W.walk(id, rev, o, k, e, found, 1u << 22)) return VGA_ERR_UNSUPPORTED;
        if (!rev && !(e & vga::POS_REV_BIT) && (e & ~vga::POS_REV_BIT) - p != k) hop++;
        const std::string* pick = nullptr;
        const size_t distinct = found.size();
        if (distinct > 1) amb++;
        if (distinct == 1) pick = &*found.begin();
        else if (hash) {
            for (const std::string& c : found)
                if (hash(c.data(), k, user) == key_hash[i]) { pick = &c; break; }
        } else if (distinct > 1) {
            // no hash function on this host: resolved below by rank (keys are stored in sorted k-mer order)
            pending.push_back(i);
            cand[{s, e}] = std::vector<std::string>(found.begin(), found.end());
            continue;
        }
        if (!pick) return VGA_ERR_INVALID;                            // no walk reproduces the key: not this index's table
        uint64_t packed = 0;
        if (!vga::pack_kmer(pick->data(), k, &packed)) return VGA_ERR_UNSUPPORTED;   // index k-mers are ACGT only (N handles are dropped)
        key_kmer2bit_out[i] = packed;
    }
    if (!pending.empty()) {
        // Rank rule.  Every k-mer that walks from `start` to `end` is a key of the index and has (start, end) in its run, and
        // nothing else has: the keys whose run holds a position pair P, in index order, are the candidate k-mers of P in sorted
        // order (the table is built from the k-mers sorted by their bytes, reference src/kmer.rs:298, 816-928).
        std::map<std::pair<uint64_t, uint64_t>, std::vector<uint64_t>> holders;
        for (auto& c : cand) holders[c.first];
        for (uint64_t i = 0; i < view->n_keys; i++) {
            for (uint64_t x = view->key_offset[i]; x < view->n_table && view->tab_start[x] != vga::POS_DELIM; x++) {
                auto it = holders.find({view->tab_start[x], view->tab_end[x]});
                if (it != holders.end() && (it->second.empty() || it->second.back() != i)) it->second.push_back(i);
            }
        }
        for (uint64_t i : pending) {
            const uint64_t off = view->key_offset[i];
            const std::pair<uint64_t, uint64_t> P{view->tab_start[off], view->tab_end[off]};
            const std::vector<uint64_t>& hs = holders[P];
            const std::vector<std::string>& cs = cand[P];
            if (hs.size() != cs.size()) return VGA_ERR_INVALID;      // the table does not come from this graph's k-mer walks
            const size_t rank = std::lower_bound(hs.begin(), hs.end(), i) - hs.begin();
            uint64_t packed = 0;
            if (rank >= cs.size() || !vga::pack_kmer(cs[rank].data(), k, &packed)) return VGA_ERR_INVALID;
            key_kmer2bit_out[i] = packed;
        }
    }
    if (n_ambiguous) *n_ambiguous = amb;
    if (n_hopping) *n_hopping = hop;
    return 0;
}
