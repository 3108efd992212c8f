// Host-side index data model for the B200 `map --also-align` path.
//
// This is the host mirror of the reference's `Index` (reference src/index.rs:36-90) in the
// flattened form the C ABI (include/vga.h, VgaIndexView) borrows.  The Rust front end would fill
// a VgaIndexView straight from its deserialised `.idx`; in this repository (no cargo in the
// image) the same arrays are produced by index_build.cpp, which restates `Index::build`
// (reference src/index.rs:109-281).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace vga {

// Packed handle, identical to handlegraph's Handle(u64): id*2 + is_reverse
// (used as such at reference src/index.rs:419-420).
typedef uint64_t handle_t;
static inline handle_t make_handle(uint64_t id, bool rev) { return id * 2 + (rev ? 1 : 0); }
static inline uint64_t handle_id(handle_t h) { return h >> 1; }
static inline bool handle_rev(handle_t h) { return h & 1; }
static inline handle_t handle_flip(handle_t h) { return h ^ 1; }

// Orientation flag of a SeqPos is carried in bit 63 of the position word
// (SeqOrient::Forward=0 < Reverse=1, reference src/kmer.rs:18-31).
static const uint64_t POS_REV_BIT = 1ull << 63;
// KMER_POS_DELIMITER {Reverse,u64::MAX,Reverse,u64::MAX} (reference src/kmer.rs:740-749).
static const uint64_t POS_DELIM = ~0ull;

// HashGraph as handlegraph 0.5.0 builds it from a GFA: per node the sequence plus the
// left/right adjacency vectors in L-line order.
struct HashGraph {
    uint64_t n_nodes = 0;                         // ids are 1..n_nodes
    std::vector<std::string> seq;                 // seq[id]
    std::vector<std::vector<handle_t>> left, right;  // left[id], right[id]
    // handle_edges_iter(h, Right) / (h, Left)
    std::vector<handle_t> edges_right(handle_t h) const;
    std::vector<handle_t> edges_left(handle_t h) const;
    std::string sequence(handle_t h) const;       // reverse handles → reverse complement
    void create_edge(handle_t l, handle_t r);
};

struct FlatIndex {
    uint32_t k = 0;
    uint64_t seq_len = 0, n_nodes = 0, n_edges = 0;
    std::string seq_fwd, seq_rev;
    std::vector<uint64_t> node_ref;       // 3*(n_nodes+1): seq_idx, edge_idx, edges_to_node
    std::vector<uint64_t> edges;          // packed handles, per node incoming then outgoing
    // k-mer table: one key per distinct k-mer in sorted (byte) order.
    std::vector<std::string> key_kmer;    // the k-mer strings (what the reference hashes with ahash)
    std::vector<uint64_t> key_kmer2bit;   // 2-bit packed (A0 C1 G2 T3, first base most significant)
    std::vector<uint64_t> key_offset;     // start index into tab_* (== bhpf.get(hash), src/index.rs:320)
    std::vector<uint64_t> tab_start, tab_end;  // kmer_pos_table incl. one POS_DELIM pair per key
    uint64_t n_kmer_pos = 0;              // entries incl. delimiters (reference n_kmer_pos)

    uint64_t seq_idx(uint64_t i) const { return node_ref[3 * i]; }
    uint64_t edge_idx(uint64_t i) const { return node_ref[3 * i + 1]; }
    uint64_t edges_in(uint64_t i) const { return node_ref[3 * i + 2]; }
};

// Parse a GFA file the way gfa 0.8.0 GFAParser + HashGraph::from_gfa do
// (reference src/subcommands/index_main.rs:72-74).  Throws std::runtime_error.
HashGraph graph_from_gfa_file(const std::string& path);
HashGraph graph_from_gfa_text(const std::string& text);

// Index::build (reference src/index.rs:109-281) without sampling.
FlatIndex build_index(const HashGraph& g, uint32_t k, uint64_t max_furcations = 100,
                      uint64_t max_degree = 100);

// Own container format (".vgidx"); the reference's bincode `.idx` needs the Rust side.
void save_index(const FlatIndex& ix, const std::string& path);
FlatIndex load_index(const std::string& path);

// The reference's own index file (bincode of `Index`, reference src/index.rs:284-305): idx_file.cpp
FlatIndex load_idx(const std::string& path);
void save_idx(const FlatIndex& ix, const std::string& path, const std::vector<uint64_t>& key_hash);

std::string reverse_complement(const std::string& s);  // reference src/dna.rs:5-40
// 2-bit packing; returns false when the k-mer has a byte outside uppercase ACGT.
bool pack_kmer(const char* s, uint32_t k, uint64_t* out);

}  // namespace vga
