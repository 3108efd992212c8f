/* vga.h — C ABI of libvgaligner_gpu.so: the B200 replacement for the body of rs-vgaligner's
 * `map_reads` loops (reference src/map.rs:56-111 and src/map.rs:162-167).
 *
 * The Rust front end keeps the CLI, `.idx` loading, FASTA/FASTQ/GFA readers and GAF writers and
 * calls these entry points through `extern "C"` (binding shown in INTEGRATION.md).  Plain pointers
 * and sizes only; no torch / C++ types.  Every function returns 0 on success and a negative
 * VGA_ERR_* code on failure (the reference panics instead: src/map.rs:141-143,
 * src/serialization.rs:52-53); vga_last_error() returns a thread-local message.  Nothing unwinds
 * across this boundary.  There is no CPU fallback: without a usable CUDA device every compute
 * entry point fails with VGA_ERR_NO_DEVICE.
 *
 * Threads: the device chosen by vga_init() belongs to the calling host thread (one thread or process per GPU).  A
 * vga_index_t carries its own workspace (streams, trace pool): calls on ONE index are serialised by the library, calls on
 * DIFFERENT indexes (also two uploads of the same index on one GPU) may run concurrently from different threads.  Result
 * objects are independent of each other and of later calls.
 */
#ifndef VGA_H
#define VGA_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define VGA_OK 0
#define VGA_ERR_INVALID (-1)      /* bad argument */
#define VGA_ERR_NO_DEVICE (-2)    /* no CUDA device / driver */
#define VGA_ERR_CUDA (-3)         /* CUDA runtime error (message has the detail) */
#define VGA_ERR_UNSUPPORTED (-4)  /* index feature outside this path (k>32, positions >= 2^31, ...) */
#define VGA_ERR_NOMEM (-5)

/* Orientation flag of a position word: bit 63 set = SeqOrient::Reverse (reference src/kmer.rs:18-31) */
#define VGA_POS_REV (1ull << 63)
/* KMER_POS_DELIMITER as a (start,end) pair of position words (reference src/kmer.rs:740-749) */
#define VGA_POS_DELIM (~0ull)

/* Borrowed view of the reference's `Index` (src/index.rs:36-90); valid only during vga_index_upload.
 * Replaces what `Index::load_from_file` (src/index.rs:296-305) hands to map_reads. */
typedef struct {
    uint32_t k;                  /* Index.kmer_length */
    uint64_t seq_len;            /* Index.seq_length */
    uint64_t n_nodes;            /* Index.n_nodes */
    uint64_t n_edges;            /* Index.n_edges */
    uint64_t n_keys;             /* Index.n_kmers */
    uint64_t n_table;            /* Index.n_kmer_pos: kmer_pos_table entries incl. delimiters */
    const uint8_t* seq_fwd;      /* seq_len ASCII bytes (Index.seq_fwd, src/index.rs:45) */
    const uint64_t* node_ref;    /* 3*(n_nodes+1): NodeRef{seq_idx,edge_idx,edges_to_node} (src/utils.rs:14-22) */
    const uint64_t* edges;       /* n_edges packed handles id*2+rev (Index.edges, src/index.rs:55) */
    const uint64_t* key_kmer2bit;/* n_keys k-mers, 2 bits/base A0 C1 G2 T3, first base most significant.
                                    Stands for the ahash keys in kmer_pos_ref (src/index.rs:74): the host
                                    re-derives each key's k-mer from its first table position. */
    const uint64_t* key_offset;  /* n_keys: bhpf.get(hash) = start index into the table (src/index.rs:320) */
    const uint64_t* tab_start;   /* n_table: KmerPos.start position | VGA_POS_REV; VGA_POS_DELIM delimiters */
    const uint64_t* tab_end;     /* n_table: KmerPos.end likewise (src/kmer.rs:732-749) */
} VgaIndexView;

/* map_reads' tunables (src/map.rs:27-44; defaults src/subcommands/map_main.rs:30-52,103) plus the
 * abPOA parameters the reference leaves at rs-abPOA's defaults (src/align.rs:202). */
typedef struct {
    uint32_t bandwidth;    /* 50 */
    uint32_t max_gap;      /* 1000 */
    uint32_t min_anchors;  /* 3 */
    uint32_t best_n;       /* 1 */
    int32_t match, mismatch, o1, e1, o2, e2; /* 2 4 4 2 24 1 */
    int32_t wb;            /* 10 */
    float wf;              /* 0.01 */
    uint32_t flags;        /* VGA_FLAG_* */
} VgaParams;

#define VGA_FLAG_NO_ALIGN 1u     /* `map` without --also-align: chains only (src/map.rs:154) */
#define VGA_FLAG_KEEP_DEBUG 2u   /* also return anchors / sorted order / f / pred / subgraphs (tests) */
#define VGA_FLAG_CHAIN_SCORES 4u /* also return ca_id / ca_score (not needed by the GAF writers) */
/* POA band granularity: the band of every DP row is widened to multiples of 8 / 16 columns, as upstream abPOA's SIMD vectors do
 * (8 = int16 lanes of SSE4.1 ... 16 = AVX2); default 1.  A compatibility switch for comparisons with a real abPOA build. */
/* Anchors of every orientation (`anchors_for_query(.., only_forward = false)`, reference src/chain.rs:154-156): reverse-strand
 * positions become anchors, are sorted Reverse-first (src/chain.rs:386-389) and chain only with anchors of the same orientation
 * (src/chain.rs:277-287).  Needs an index uploaded with VGA_INDEX_ALL_ORIENTATIONS; ca_tb / ca_te carry the orientation in
 * bit 31.  Chains with a reverse-strand anchor are not aligned (aln_status = VGA_ALN_UNSUPPORTED): the reference's Reverse / Both
 * ranges subtract forward node starts from reverse-strand positions (src/align.rs:536-547 with src/index.rs:461-480), which is
 * not a defined computation. */
#define VGA_FLAG_ALL_ORIENTATIONS 0x1000u
#define VGA_FLAG_BAND_GRAN8 0x100u
#define VGA_FLAG_BAND_GRAN16 0x200u

/* per-read alignment status (aln_status) */
#define VGA_ALN_PLACEHOLDER 0    /* no chain: GAFAlignment::from_placeholder_chain (src/align.rs:913-930) */
#define VGA_ALN_OK 1
#define VGA_ALN_EMPTY 2          /* no end row reached the last query column */
#define VGA_ALN_TRACE_ERR 3
#define VGA_ALN_BAND_ERR 4
#define VGA_ALN_UNSUPPORTED 5    /* Reverse/Both node range (src/align.rs:365-387), unreachable from map_reads */
#define VGA_ALN_TOO_WIDE 6       /* adaptive band wider than 1024 columns on a read too long for the shared-memory row of the wide kernel (~38 k bases) */

/* Flat, library-owned (pinned host) result arrays for one batch of n reads; valid until
 * vga_results_free().  A read with no chain (the reference's placeholder chain, src/chain.rs:644-649)
 * has zero chains here. */
typedef struct {
    uint32_t n_reads;
    /* per read */
    const uint32_t* n_anchors;        /* n: anchors_for_query(..., only_forward=true) count (src/map.rs:62) */
    const uint64_t* chain_off;        /* n+1: chain index range of each read */
    /* per chain */
    uint64_t n_chains;
    const uint64_t* chain_anchor_off; /* n_chains+1 */
    /* per chain anchor, in chain (query) order (Chain.anchors, src/chain.rs:176-187) */
    uint64_t n_chain_anchors;
    const uint32_t* ca_id;            /* Anchor.id (VGA_FLAG_CHAIN_SCORES or KEEP_DEBUG, else NULL) */
    const uint32_t* ca_qb;            /* Anchor.query_begin (query_end = +k) */
    const uint32_t* ca_tb;            /* Anchor.target_begin.position */
    const uint32_t* ca_te;            /* Anchor.target_end.position */
    const double* ca_score;           /* Anchor.max_chain_score (same flags, else NULL) */
    /* best alignment per read (best_alignment_for_query, src/align.rs:34-55) */
    const uint8_t* aln_status;        /* n: VGA_ALN_* */
    const uint32_t* aln_chain;        /* n: index (within the read) of the chain that was kept */
    const uint64_t* aln_path_off;     /* n+1 */
    const uint64_t* aln_path;         /* packed handles: graph_nodes.dedup() mapped through the range (src/align.rs:1107-1123) */
    const uint64_t* aln_run_off;      /* n+1 */
    const uint32_t* aln_runs;         /* op<<28 | len; op 0 '=' 1 'X' 2 'D' 3 'I' (-> cs string and cigar) */
    const uint32_t* aln_path_len;     /* n: |abpoa_nodes| (src/align.rs:1152) */
    const uint32_t* aln_start_off;    /* n: aln_start_offset (src/align.rs:1155) */
    const uint32_t* aln_end_off;      /* n: aln_end_offset (src/align.rs:1156) */
    const uint32_t* aln_n_aligned;    /* n: n_aligned_bases (src/align.rs:1158) */
    const int32_t* aln_score;         /* n: POA score (not printed by the reference) */
    const uint64_t* aln_cells;        /* n: DP cells computed for the kept alignment (GCUPS numerator) */
    /* VGA_FLAG_KEEP_DEBUG only, else NULL */
    const uint64_t* dbg_anchor_off;   /* n+1 */
    const uint32_t* dbg_qb; const uint32_t* dbg_tb; const uint32_t* dbg_te;  /* anchors in emission order */
    const uint32_t* dbg_sorted;       /* anchor id at each sorted position (src/chain.rs:386-389) */
    const double* dbg_f;              /* max_chain_score by sorted position */
    const int32_t* dbg_pred;          /* best_predecessor_id by sorted position, -1 = None */
    const uint64_t* dbg_sub_off;      /* n+1 */
    const uint64_t* dbg_sub_handles;  /* extended node range handed to the aligner (src/align.rs:82) */
    const uint64_t* dbg_edge_off;     /* n+1 (edge index) */
    const uint32_t* dbg_edges;        /* (s,e) pairs (src/align.rs:670-724) */
} VgaResultsView;

typedef struct {
    uint64_t n_reads, n_bases, n_probes, n_positions_read, n_anchors, n_chains, n_alignments;
    uint64_t poa_cells;                     /* all aligned (read,chain) pairs */
    uint64_t lookup_bytes;                  /* algorithmic bytes of the lookup stage */
    float ms_h2d, ms_seed, ms_sort, ms_chain, ms_subgraph, ms_poa, ms_trace, ms_d2h;
    float ms_total;                         /* CUDA-event time of the whole call on the library's stream */
    float ms_wall;                          /* host wall clock of the call */
    uint32_t kernel_launches;               /* kernels of this library launched by the call */
    uint32_t poa_launches;                  /* launches of the POA fill kernel among them */
    uint64_t h2d_bytes, d2h_bytes;          /* bytes copied host->device / device->host by the call */
} VgaStats;

typedef struct vga_index vga_index_t;
typedef struct vga_results vga_results_t;

/* Select the CUDA device this process drives (one process per GPU; under torchrun pass LOCAL_RANK).
 * Compute entry points called before vga_init() use device 0. */
int vga_init(int device);
int vga_device_count(void);
/* Build the device-resident k-mer -> position table and graph arrays (replaces the use of the loaded
 * `Index` inside map_reads, reference src/map.rs:27-29).  The image is ONE device allocation with
 * internal offsets, so replicating it to the other GPUs of the box is a single NCCL broadcast:
 * rank 0 uploads, copies the image into a byte tensor (vga_index_image_copy_to_device), broadcasts,
 * and every other rank adopts its copy with vga_index_import_device. */
int vga_index_upload(const VgaIndexView* view, vga_index_t** out);
/* flags: VGA_INDEX_ALL_ORIENTATIONS keeps every table position (orientation in bit 31 of the device words) instead of the
 * forward/forward ones `map_reads` uses (src/map.rs:62); such an index serves vga_map_batch only with VGA_FLAG_ALL_ORIENTATIONS. */
#define VGA_INDEX_ALL_ORIENTATIONS 1u
int vga_index_upload_ex(const VgaIndexView* view, uint32_t flags, vga_index_t** out);
int vga_index_free(vga_index_t* ix);
int vga_index_image_size(const vga_index_t* ix, uint64_t* bytes);
int vga_index_image_copy_to_device(const vga_index_t* ix, void* dev_dst);
int vga_index_import_device(const void* dev_blob, uint64_t bytes, vga_index_t** out);

/* Multi-GPU behind the ABI (no torch on a Rust host): one process or host thread per GPU.  Rank 0 creates an id
 * (ncclGetUniqueId) and hands its VGA_COMM_ID_BYTES bytes to the other ranks by any means; then EVERY rank calls
 * vga_index_broadcast after vga_init(its device): rank 0 with its uploaded index, the others with *ix = NULL, which receive a
 * replica — ONE ncclBroadcast of the index image, the only collective of the path.  Each rank then maps its contiguous range
 * of reads with vga_map_batch; outputs are concatenated in rank order.  NCCL is loaded at run time (libnccl.so.2). */
#define VGA_COMM_ID_BYTES 128
int vga_comm_unique_id(void* id_bytes);
int vga_index_broadcast(vga_index_t** ix, const void* id_bytes, int rank, int world);

/* GPU index build (reference src/index.rs:109-281): k-mer walks (src/kmer.rs:347-505), stable sort + consecutive dedup
 * (:298-301) and the position table (:816-928) on the device.  The host passes the linearised graph (seq_fwd, NodeRef, edges —
 * the first fields of VgaIndexView; src/utils.rs:81-146 stays on the host) and gets the table arrays of VgaIndexView back,
 * identical to what the reference's build produces (defaults max_furcations = max_degree = 100).  Graphs whose walks explode
 * (cyclic, out-degree in the tens) or that reach the reference's degree limits return VGA_ERR_UNSUPPORTED: build those on the host. */
typedef struct {
    uint64_t seq_len, n_nodes, n_edges;
    const uint8_t* seq_fwd;
    const uint64_t* node_ref;   /* 3 * (n_nodes + 1) */
    const uint64_t* edges;
} VgaGraphView;
typedef struct vga_index_build vga_index_build_t;
int vga_index_build_device(const VgaGraphView* graph, uint32_t k, vga_index_build_t** out);
/* fills `view` with the graph's pointers and the built table (borrowed from `b`); *ms_device = device time of the build */
int vga_index_build_view(const vga_index_build_t* b, const VgaGraphView* graph, uint32_t k, VgaIndexView* view, float* ms_device);
int vga_index_build_free(vga_index_build_t* b);

void vga_default_params(VgaParams* p);
/* Map one batch: reads as concatenated bytes + n_reads+1 offsets (host memory).  Blocking. */
int vga_map_batch(vga_index_t* ix, const char* seqs, const uint64_t* seq_off, uint32_t n_reads,
                  const VgaParams* params, vga_results_t** out);
/* Same with the reads already resident on the index's device 0 (device-timed benchmark leg). */
int vga_map_batch_device(vga_index_t* ix, const char* dev_seqs, const uint64_t* dev_seq_off, uint32_t n_reads,
                         uint64_t n_bases, const VgaParams* params, vga_results_t** out);
/* Anchors per read (the K1 counting stage; equal to VgaResultsView.n_anchors of vga_map_batch) WITHOUT any device allocation.
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Reads and offsets are resident on the index's device, and so
 * is the output dev_n_anchors[n_reads]; n_bases = dev_seq_off[n_reads] - dev_seq_off[0].  The stage's scratch is a fixed device
 * array of VGA_NOALLOC_SCRATCH_BYTES inside the library (held by every GPU that loads it, until exit): a call needs
 * 8 * (n_reads + 1) + 8 * n_bases bytes of it and is rejected with VGA_ERR_INVALID beyond that.  Every call shares the array:
 * calls that overlap on one GPU race (calls of one process are serialised by the library).  Blocking. */
#define VGA_NOALLOC_SCRATCH_BYTES (64u << 20)
int vga_seed_count_noalloc(vga_index_t* ix, const char* dev_seqs, const uint64_t* dev_seq_off, uint32_t n_reads,
                           uint64_t n_bases, uint32_t* dev_n_anchors);
int vga_results_view(const vga_results_t* r, VgaResultsView* view);
int vga_results_stats(const vga_results_t* r, VgaStats* stats);
int vga_results_free(vga_results_t* r);
const char* vga_last_error(void);

/* measured INT32 ALU throughput of the current device (ops/s; IADD3/VIMNMX streams) for the POA roofline */
int vga_measure_int32_peak(double* ops_per_s);

#ifdef __cplusplus
}
#endif
#endif
