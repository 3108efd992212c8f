/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU oracle for the `vgaligner map --also-align` hot path (reference src/map.rs:27-216).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product path (rs-vgaligner_b200/csrc, libvgaligner_gpu.so) never does.
 *
 * Parity status:
 *   seeding, index accessors, chaining, range/extension/subgraph, GAF records:
 *       restated from the reference sources (citations in oracle.cpp) and PINNED against the
 *       reference's own known-answer tests (tests/test_oracle_golden.py).
 *   POA (align_with_POA_aligner -> ab_poa::AbpoaAligner::create_align_safe, src/align.rs:202):
 *       PARITY UNPINNED.  The arithmetic lives in crate ab_poa 1.5.0 (git HopedWall/rs-abpoa, no
 *       revision pinned in Cargo.toml:38 / Cargo.lock:5-15) wrapping the C library abPOA; neither
 *       is under /root/reference and the reference has no test that pins a result at that call.
 *       oracle.cpp restates abPOA's published algorithm (adaptive band, convex gap, value-driven
 *       backtrack) and DEFINES every tie-break; see DESIGN.md "POA definition".
 */
#ifndef VGA_ORACLE_H
#define VGA_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_index orc_index_t;
typedef struct orc_results orc_results_t;

typedef struct {
    uint32_t bandwidth;    /* chaining predecessor window, 50 (src/subcommands/map_main.rs:103) */
    uint32_t max_gap;      /* 1000 (map_main.rs:30-34) */
    uint32_t min_anchors;  /* 3 (map_main.rs:42-46) */
    uint32_t best_n;       /* 1 (map_main.rs:48-52) */
    int32_t match, mismatch, o1, e1, o2, e2;  /* abPOA defaults 2,4,4,2,24,1 */
    int32_t wb;            /* 10 */
    float wf;              /* 0.01 */
    uint32_t flags;        /* ORC_FLAG_* */
} orc_params_t;

#define ORC_FLAG_NO_ALIGN 1u        /* chains only (map without --also-align) */
#define ORC_FLAG_KEEP_DEBUG 2u      /* keep per-read anchors / sorted order / f / pred / subgraphs */
#define ORC_FLAG_LITERAL_EXTEND 4u  /* extend_range_chain_2 exactly as written (no visited set) */
/* abPOA-compatibility switches of the POA definition.  Real abPOA could not be run here, so each
 * choice it might make differently is a switch that flips a constant, not a kernel: */
#define ORC_FLAG_ALL_ORIENTATIONS 0x1000u /* anchors_for_query(.., only_forward = false), src/chain.rs:154-156; positions carry the
                                            orientation in bit 31 of ca_tb / ca_te; chains with a reverse anchor are not aligned */
#define ORC_FLAG_BAND_GRAN8 0x100u  /* band of every row widened to multiples of 8 columns (SIMD vector of 8 int16 lanes) ... */
#define ORC_FLAG_BAND_GRAN16 0x200u /* ... or of 16 columns (AVX2); same bits as VGA_FLAG_BAND_GRAN8 / 16, the CUDA path follows */
#define ORC_FLAG_INT16_SCORES 0x400u /* -inf is INT16_MIN + the largest single-step penalty (upstream's int16 score width) instead of
                                        -2^29: only differs from the default when a reachable score comes near -32768 */
#define ORC_FLAG_NO_CS_PREFIX 0x800u /* alignments GAF: the cs string without its "cs:Z:" tag (rs-abPOA's format is [RECALLED]) */

void orc_default_params(orc_params_t* p);

/* Index arrays as the reference's `Index` holds them (src/index.rs:36-90).  keys = n_keys*k bytes
 * (the k-mer strings the reference hashes), key_offset = start offsets into the table, table =
 * kmer_pos_table with delimiter entries, orientation in bit 63. Everything is copied. */
orc_index_t* orc_index_create(uint32_t k, uint64_t seq_len, uint64_t n_nodes, uint64_t n_edges,
                              const char* seq_fwd, const uint64_t* node_ref, const uint64_t* edges,
                              uint64_t n_keys, const char* keys, const uint64_t* key_offset,
                              uint64_t n_tab, const uint64_t* tab_start, const uint64_t* tab_end);
void orc_index_free(orc_index_t*);

/* index accessors pinned by the reference tests (src/index.rs:388-606) */
uint64_t orc_bv_rank(const orc_index_t*, uint64_t pos);
uint64_t orc_bv_inverse_rank(const orc_index_t*, uint64_t pos);
uint64_t orc_bv_select(const orc_index_t*, uint64_t element_no);
uint64_t orc_handle_from_seqpos(const orc_index_t*, int reverse, uint64_t pos);
/* writes up to cap handles, returns the count */
uint64_t orc_incoming(const orc_index_t*, uint64_t handle, uint64_t* out, uint64_t cap);
uint64_t orc_outgoing(const orc_index_t*, uint64_t handle, uint64_t* out, uint64_t cap);
uint64_t orc_seq_from_handle(const orc_index_t*, uint64_t handle, char* out, uint64_t cap);
/* find_positions_for_query_kmer (src/index.rs:353-382): returns count; start/end carry bit 63 */
uint64_t orc_positions(const orc_index_t*, const char* kmer, uint64_t len, uint64_t* start, uint64_t* end,
                       uint64_t cap);
/* anchors_for_query (src/chain.rs:134-173) */
uint64_t orc_anchors(const orc_index_t*, const char* read, uint64_t len, int only_forward, uint64_t* qb,
                     uint64_t* qe, uint64_t* tb, uint64_t* te, uint64_t cap);
/* score_anchor (src/chain.rs:274-368); orientation flags 0 = Forward */
double orc_score_anchor(uint64_t a_qb, uint64_t a_qe, uint64_t a_tb, uint64_t a_te, double a_f, uint64_t b_qb,
                        uint64_t b_qe, uint64_t b_tb, uint64_t b_te, uint64_t seed_length, uint64_t max_gap);

/* map_reads (src/map.rs:27-216) over reads given as concatenated bytes + n+1 offsets.
 * n_threads > 1 parallelises over reads (std::thread workers) for the CPU baseline; results are identical. */
orc_results_t* orc_map_reads(const orc_index_t*, const char* seqs, const uint64_t* seq_off, uint32_t n_reads,
                             const orc_params_t*, int n_threads);
void orc_results_free(orc_results_t*);
/* flat result arrays by name; returns pointer and element count (NULL if unknown) */
const void* orc_results_field(const orc_results_t*, const char* name, uint64_t* count);
/* GAF text (src/align.rs:971-1027); names = '\n'-separated read names; which: 0 chains, 1 alignments */
const char* orc_results_gaf(orc_results_t*, const orc_index_t*, const char* names, const char* seqs,
                            const uint64_t* seq_off, int which, uint64_t* len);

/* stand-alone POA (the FFI seam at src/align.rs:190-202): labels concatenated + n+1 offsets, edges
 * as (s,e) pairs; fills one-read results.  Used by unit tests. */
orc_results_t* orc_poa_align(const char* labels, const uint64_t* label_off, uint32_t n_nodes,
                             const uint32_t* edges, uint32_t n_edges, const char* query, uint32_t qlen,
                             const orc_params_t*);

#ifdef __cplusplus
}
#endif
#endif
