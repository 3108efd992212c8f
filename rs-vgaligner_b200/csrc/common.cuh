// Shared declarations for libvgaligner_gpu.so (sm_100a).  Product code: never includes oracle/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/vga.h"

typedef uint64_t u64;
typedef uint32_t u32;
typedef int32_t i32;
typedef int64_t i64;
typedef uint16_t u16;
typedef uint8_t u8;

namespace vga {

void set_error(const std::string& msg);
const char* get_error();

struct CudaFail { int code; };   // thrown inside the library, caught at the C boundary

#define VGA_CUDA(call)                                                                              \
    do {                                                                                            \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess) {                                                                   \
            vga::set_error(std::string(#call) + ": " + cudaGetErrorString(e__) + " (" + __FILE__ + ":" + \
                           std::to_string(__LINE__) + ")");                                         \
            throw vga::CudaFail{e__ == cudaErrorMemoryAllocation ? VGA_ERR_NOMEM : VGA_ERR_CUDA};   \
        }                                                                                           \
    } while (0)

// ---------------------------------------------------------------------------------------------
// Device-resident index image: ONE allocation, position independent (offsets from the blob base),
// so that a single NCCL broadcast replicates it (BASELINE.json north_star).
// ---------------------------------------------------------------------------------------------
struct Slot {          // open-addressed k-mer table entry, 16 B (two per 32-B DRAM sector)
    u64 key;           // 2-bit packed k-mer
    u32 off;           // first position in pos_tb/pos_te
    u32 cnt;           // number of forward/forward positions; 0 = empty slot
};

struct IndexHeader {
    u64 magic;
    u64 bytes;          // whole image
    u32 k, n_nodes, n_edges, L;
    u64 n_keys, n_pos, n_slots;   // n_slots is a power of two
    u64 off_seq_both;   // u8[2L]: seq_fwd then seq_rev (Index.seq_fwd / seq_rev, src/index.rs:45-47)
    u64 off_seq_idx;    // u32[n_nodes+1]  NodeRef.seq_idx
    u64 off_edge_idx;   // u32[n_nodes+1]  NodeRef.edge_idx
    u64 off_edges_in;   // u32[n_nodes+1]  NodeRef.edges_to_node
    u64 off_edges;      // u32[n_edges]    packed handles
    u64 off_pos2node;   // u32[L]          rank1(seq_bv, pos) = node id of a forward position
    u64 off_slots;      // Slot[n_slots]
    u64 off_pos_tb;     // u32[n_pos]      KmerPos.start.position of fwd/fwd entries, table order
    u64 off_pos_te;     // u32[n_pos]      KmerPos.end.position
    // compact table (k <= 16; see below): 64-byte buckets of 8 entries and the interleaved position array; n_buckets = 0: absent
    u64 off_buckets;    // uint4[4 * n_buckets]
    u64 n_buckets;
    u64 off_pos2;       // uint2[n_pos2]   {start, end} of fwd/fwd entries, table order; a run with more than 30 entries is
    u64 n_pos2;         //                 preceded by a header element {count, 0}
    u64 all_orient;     // 1: the table holds positions of every orientation (bit 31 of pos_tb / pos_te = Reverse)
    u64 pad[3];
};
static const u64 INDEX_MAGIC = 0x3358444947414756ull;  // "VGAGIDX3" (slot keys in bit-plane form; compact bucket table)

// Compact k-mer table for k <= 16 (both bit planes of a key fit 16 bits each): 8-byte entries in 64-byte buckets — one
// DRAM access (64 B on HBM3e) answers a probe, where a 16-byte slot of the open-addressed table cost 1.4 accesses of
// 64 B for 16 useful bytes — and a table a quarter of the size (C4: 134 MB instead of 537, most of it L2-resident).
//   entry = key32 | payload << 32;  key32 = plane0 | plane1 << 16;  an all-zero entry is empty
//   payload, one fwd/fwd position (the usual case): bit 31 set, start in bits 0..25, end - start - k in bits 26..30:
//            the anchor comes out of the probe itself, the position array is not touched
//   payload otherwise: first element in pos2 in bits 0..25, count in bits 26..30 (31 = the run has a header element)
// A key lives in the first bucket from its home bucket on that has room; a probe walks buckets until it finds the key or a
// bucket with an empty last entry.  Entries of a bucket are filled front to back and never removed.
static const u32 CT_INLINE = 0x80000000u, CT_LOW26 = 0x03ffffffu;
static const u32 CT_MAX_INLINE_DLT = 31, CT_MAX_CNT = 30;
__host__ __device__ static inline u32 ct_bucket(u32 key32, u32 n_buckets) {
    u32 h = key32 * 0x9E3779B1u;
    h ^= h >> 15;
    h *= 0x85EBCA77u;
    h ^= h >> 13;
#ifdef __CUDA_ARCH__
    return __umulhi(h, n_buckets);
#else
    return (u32)(((u64)h * n_buckets) >> 32);
#endif
}

struct DevIndex {       // resolved device pointers, passed to kernels by value
    u32 k, n_nodes, n_edges, L;
    u32 slot_mask;
    const u8* seq_both;
    const u32* seq_idx;
    const u32* edge_idx;
    const u32* edges_in;
    const u32* edges;
    const u32* pos2node;
    const Slot* slots;
    const u32* pos_tb;
    const u32* pos_te;
    const uint4* buckets;   // compact table (n_buckets > 0): the seeding kernels use it instead of slots / pos_tb / pos_te
    const uint2* pos2;
    u32 n_buckets;
};

// Slot keys are stored in BIT-PLANE form: bit t of the low word = bit 0 of the 2-bit code (A0 C1 G2 T3) of base t of
// the k-mer (t = 0 is the first base), bit t of the high word = bit 1.  The seeding kernel gets both words of a
// query k-mer from two warp ballots and a funnel shift, without packing bases one by one.
__host__ __device__ static inline u64 kmer_planes_from_2bit(u64 key2bit, u32 k) {
    u64 lo = 0, hi = 0;
    for (u32 t = 0; t < k; t++) {
        const u64 c = (key2bit >> (2 * (k - 1 - t))) & 3u;     // first base is the most significant pair (VgaIndexView.key_kmer2bit)
        lo |= (c & 1u) << t;
        hi |= (c >> 1) << t;
    }
    return (hi << 32) | lo;
}
__host__ __device__ static inline u64 kmer_slot_hash(u64 key) {
    key *= 0x9E3779B97F4A7C15ull;
    key ^= key >> 29;
    key *= 0xBF58476D1CE4E5B9ull;
    key ^= key >> 32;
    return key;
}

// ---------------------------------------------------------------------------------------------
// POA constants shared by fill and traceback kernels
// ---------------------------------------------------------------------------------------------
static const i32 POA_NEG_INF = -(1 << 29);   // same sentinel as the oracle
static const i32 POA_DEEP = -(1 << 30);      // generator value of masked cells (never wins a max)

// Band granularity (VGA_FLAG_BAND_GRAN8 / 16): upstream abPOA computes whole SIMD vectors, so its effective band
// begins and ends on vector boundaries.  After the band of the row is known (adaptive rule + clipping to the predecessors), it is
// widened to multiples of g = gmask + 1 columns and clipped again on the right; gmask = 0 leaves it as it is.  Every band begin is
// then a multiple of g (the source row begins at 0), so the widened begin never leaves the predecessors' bands.
__host__ __device__ static inline void band_round(i32& beg, i32& end, i32 gmask, i32 qlen, i32 max_pe) {
    beg &= ~gmask;
    i32 e = end | gmask;
    if (e > qlen) e = qlen;
    if (e > max_pe + 1) e = max_pe + 1;
    end = e;
}
__host__ __device__ static inline i32 band_round_end0(i32 end0, i32 gmask, i32 qlen) { const i32 e = end0 | gmask; return e > qlen ? qlen : e; }

// per DP row: band, trace location and the reference score the row's trace words are relative to
// toff: pool offset in 4-word (16 B) units; prow: row of the first predecessor (relative to the item's
// row_base); meta: number of predecessor rows (low 24 bits, 0 = virtual source) | graph base << 24;
// node/off: subgraph node index and offset inside its label.  Everything the traceback needs for a
// row with a single predecessor is in this one 32-byte record.
// trace row layout: the word of column j of a row with band [beg,end] is pool[toff*4 + j - trace_col0(beg)], i.e. rows are
// stored from the 8-column block that contains beg (the pair kernel stores whole blocks with 16-byte vectors)
__host__ __device__ static inline i32 trace_col0(i32 beg) { return beg & ~7; }
struct RowInfo { i32 beg, end; u32 toff; i32 ref; u32 prow, meta, node, off; };
// per subgraph node: band, row-maximum span and maximum of the node's LAST row
struct NodeState { i32 beg, end, lpos, rpos, mx, pad0, pad1, pad2; };

// per-item status while in flight (device side); final values are VGA_ALN_*
enum : u32 {
    IT_PENDING = 100,
    IT_RETRY_WIDER = 101,   // does not fit this kernel instance (band wider than 32*C columns, or too many rows for FIXED)
    IT_RETRY_POOL = 102,    // trace pool exhausted in this wave
};

struct AlignItem {      // one (read, chain) alignment job
    u32 read;           // read index within the chunk
    u32 chain;          // global chain index within the chunk
    u32 S;              // subgraph nodes
    u32 nrows;          // subgraph bases
    u32 qlen;
    u32 node_base;      // into nd_* arrays
    u32 pred_base;      // into nd_pred
    u32 row_base;       // into row_* arrays
    u64 q_off;          // into the chunk's read bytes
    u32 ext_off, ext_cnt;   // extension handles in the ext pool
    u32 min_node, max_node; // node-id range of the chain (find_range_chain)
    u32 E;              // pred entries
    u32 out_base;       // into per-item traceback output regions (runs)
    u32 status;         // IT_* / VGA_ALN_*
    i32 score;
    u32 end_node;       // node whose last row ends the alignment
    u32 n_runs, n_path, path_len, start_off, end_off, n_aligned;
    u32 run_pos, path_pos;  // where the (back-to-front) traceback output starts inside its region
    u64 cells;
};

}  // namespace vga
