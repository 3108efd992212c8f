// Internal declarations shared by the translation units of libvgaligner_gpu.so.
#pragma once
#include <mutex>
#include <vector>

#include "common.cuh"

struct vga_workspace;

struct vga_index {
    std::mutex mu;           // one call per index at a time (the workspace is per index)
    int device = 0;
    void* d_blob = nullptr;
    vga::IndexHeader hdr;
    vga::DevIndex dv;
    vga_workspace* ws = nullptr;
};

namespace vga {

extern thread_local int g_device;   // device selected by vga_init on this host thread
void free_workspace(vga_index* ix);

// ---- stage launchers (each returns the number of kernels it launched) -----------------------
struct SeedBuffers {
    const char* seqs;        // chunk read bytes (device)
    const u64* seq_off;      // n+1, absolute into seqs
    u32 n_reads;
    u64* read_koff;          // n+1: k-mer index range of each read
    u32* kcnt;               // per k-mer: matching positions
    u32* koff;               // per k-mer: first position in pos_tb/pos_te
    u32* read_acnt;          // n: anchors per read
    u32* read_aoff;          // n+1: exclusive scan of read_acnt
    // anchors in emission order (i asc, table order) — Anchor{id implicit, qb, tb, te}
    u32* a_qb; u32* a_tb; u32* a_te;
    u32* sort_key_in; u32* sort_val_in;   // te / local id
};

int launch_read_prep(const DevIndex& ix, const SeedBuffers& b, cudaStream_t st);
int launch_seed_count(const DevIndex& ix, const SeedBuffers& b, cudaStream_t st);
int launch_seed_emit(const DevIndex& ix, const SeedBuffers& b, cudaStream_t st);

static const int CHAIN_KMAX = 2048;   // gap lengths the integer DP keeps in shared memory (max_gap must be below)
static const int CHAIN_SMAX = 512;   // anchors per read the fused sort+chain kernel keeps in shared memory
struct ChainBuffers {
    u32 n_reads;
    const u32* read_aoff;
    const u32* a_qb; const u32* a_tb; const u32* a_te;
    u32* sidx;                // sorted position -> local anchor id (per read segment); written by the fused kernel
    const u32* big_list; u32 n_big;   // reads with more than CHAIN_SMAX anchors (sorted by CUB)
    const double* gapcost;    // [max_gap+1]: 0.01*k*g + 0.5*log2(g), host-computed (src/chain.rs:348-354)
    const i32* gapcost_i;     // [max_gap+1]: round(1000 * gapcost) when the all-integer DP is exact (vga_api.cu upload_gapcost), else NULL
    double* f;                // by sorted position
    i32* pred;                // sorted position of the best predecessor, -1 = None (consumed by backtracking)
    i32* pred_dbg;            // copy taken before backtracking, as anchor ids (NULL unless debug)
    u32* n_chains;            // per read
    u32* n_members;           // per read: total chain anchors
    u32* clen;                // chain lengths, region of read r at read_aoff[r] + r
    u32* cmem;                // chain members (sorted positions), region at 2*read_aoff[r] + 2*r
    u32 k, bandwidth, max_gap, min_anchors;
};
int launch_chain_prep(const u64* aoff64, u32 n, u32* aoff32, u32* big_ctr, u32* big_list, u32* big_beg, u32* big_end,
                      cudaStream_t st);
int launch_chain(const ChainBuffers& b, bool fused, cudaStream_t st);

struct ChainGather {
    u32 n_reads;
    const u32* read_aoff; const u32* a_qb; const u32* a_tb; const u32* a_te; const u32* sidx;
    const double* f;
    const u32* n_chains; const u32* clen; const u32* cmem;
    const u64* chain_off;     // n+1: exclusive scan of n_chains
    const u64* member_off;    // n+1: exclusive scan of n_members
    u64* chain_anchor_off;    // n_chains_total+1
    u32* ca_id; u32* ca_qb; u32* ca_tb; u32* ca_te; double* ca_score;
};
int launch_chain_gather(const ChainGather& g, cudaStream_t st);

struct SubBuffers {
    u32 n_items;
    AlignItem* items;
    const char* seqs; const u64* seq_off;
    const u64* chain_anchor_off; const u32* ca_qb; const u32* ca_tb; const u32* ca_te;
    // extension pool (bump allocated) + per-warp BFS scratch
    u32* ext_pool; u32 ext_cap; u32* ext_ctr; u32* overflow;
    u32* scratch; u32 scratch_per_warp;   // u32 words per warp
    u32 n_warps;
    // node arrays (filled by sub_fill at AlignItem.node_base / pred_base)
    u32* nd_handle; u32* nd_row_start; u32* nd_src; u32* nd_pred_off; u32* nd_pred; i32* nd_first_succ;
    i32* nd_remain_last;
};
int launch_sub_count(const DevIndex& ix, const SubBuffers& b, cudaStream_t st);
int launch_sub_fill(const DevIndex& ix, const SubBuffers& b, cudaStream_t st);
int launch_sub_rowmeta(const DevIndex& ix, const SubBuffers& b, u32* row_meta, cudaStream_t st);   // per-row records (k_subgraph.cu)

struct TraceBuffers;
struct PoaBuffers {
    u32 n_items;             // entries of work[]
    const u32* work;         // item indices to process in this launch
    AlignItem* items;
    const char* seqs;
    const u32* nd_row_start; const u32* nd_src; const u32* nd_pred_off; const u32* nd_pred;
    const i32* nd_first_succ; const i32* nd_remain_last;
    NodeState* nd_state;     // per node, written by the fill kernel for later nodes
    RowInfo* rows;           // per row, written by the fill kernel for later rows / traceback
    const u32* row_meta;     // per row: base | kind << 8 | last << 10 | remain << 11 (sub_rowmeta_kernel)
    // trace pool: one 32-bit word per band cell
    u32* pool; u64 pool_words; unsigned long long* pool_ctr;
    u32* work_ctr;
    i32 match, mismatch, o1, e1, o2, e2, wb; float wf;
    i32 nb_hint;             // pair kernel: blocks per lane and step worth instantiating (2 or 3), from rows per query base
    i32 long_mode;           // kernel family of class 3 for this call (poa_long_mode)
    i32 qmax_short;          // longest read among the items of the short-read classes (sizes the pair kernel's shared memory)
    i32 qmax_wide;           // longest read among the items of a wide-class launch (sizes its shared-memory row)
    i32 bgran_mask;          // band granularity - 1 (0, 7 or 15; band_round in common.cuh)
};
// kernel classes: 0..1 = short reads on the 16x2 kernel (k_poa_pair.cu; 1 = with the per-row range check); 2 = short reads, 32 lanes
// (k_poa_short.cu); 3..7 = SLIDE C=8,12,16,24,32 (k_poa.cu); 8 = bands wider than 1024 columns (k_poa_wide.cu).  VGA_POA_PAIR=0 puts the 8/16/32-lane kernels of k_poa_short.cu on 0..2.
int poa_long_mode(u64 n_long_items, int n_sms);   // long-read kernel family for the batch: 1 = shared-memory kernel, 0 = SLIDE (k_poa.cu)
void poa_begin_batch();     // re-reads the per-batch test / A-B switches of the class selection
int poa_class_count();
bool poa_pair_active();    // classes 0..1 run on the 16x2 kernel (plane trace rows); false with VGA_POA_PAIR=0
int poa_class_for(i64 qlen, i64 nrows, i64 west, const VgaParams& P, int long_mode);
int poa_class_wider(int cls, int long_mode);    // -1: none
int launch_poa_fill(const DevIndex& ix, const PoaBuffers& b, int cls, int n_sms, cudaStream_t st, const TraceBuffers* fused_tb = nullptr);
int launch_poa_fill_short(const DevIndex& ix, const PoaBuffers& b, int cls, int n_sms, cudaStream_t st);
bool poa_short_ok(i64 qlen, i64 nrows, const VgaParams& P);   // read short enough and scores within int16
int launch_poa_fill_long(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st, const TraceBuffers* fused_tb = nullptr);   // k_poa_long.cu
int launch_poa_fill_wide(const DevIndex& ix, const PoaBuffers& b, int n_sms, cudaStream_t st);              // k_poa_wide.cu
int launch_poa_fill_pair(const DevIndex& ix, const PoaBuffers& b, int cls, int n_sms, cudaStream_t st);   // k_poa_pair.cu
bool poa_pair_checked_ok(i64 qlen, i64 nrows, const VgaParams& P);   // same, scores checked per row instead of bounded a priori
bool poa_pair_ok(i64 qlen, i64 nrows, const VgaParams& P);    // 16x2 kernel: short read, int16 scores, default-sized gap model

struct TraceBuffers {
    u32 n_items; const u32* work; AlignItem* items;
    const char* seqs;
    const u32* nd_row_start; const u32* nd_src; const u32* nd_pred_off; const u32* nd_pred; const i32* nd_remain_last;
    const RowInfo* rows; const u32* pool;
    u32* out_runs;           // per item region at out_base, capacity nrows+qlen+2, written back to front
    u32* out_path;           // per item region at node_base, capacity S, written back to front
    i32 match, mismatch, o1, e1, o2, e2, wb; float wf;
    i32 bgran_mask;
};
int launch_poa_trace(const DevIndex& ix, const TraceBuffers& b, cudaStream_t st);

double measure_int32_peak();

}  // namespace vga
