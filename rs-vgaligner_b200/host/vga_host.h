/* C API of libvga_host.so: the host-side pieces the Rust front end owns in the real deployment
 * (index build, FASTA/FASTQ reading, GAF writing), restated in C++ because this image has no cargo.
 * Pure CPU code, no CUDA, never touches oracle/.  Used by the vga_map harness, bench.py and tests. */
#ifndef VGA_HOST_H
#define VGA_HOST_H
#include <stdint.h>

#include "../../include/vga.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct vgah_index vgah_index_t;
typedef struct vgah_reads vgah_reads_t;

const char* vgah_last_error(void);

/* `vgaligner index` (reference src/subcommands/index_main.rs:11-87, src/index.rs:109-281) */
vgah_index_t* vgah_index_build_file(const char* gfa_path, uint32_t k, uint64_t max_furcations, uint64_t max_degree);
vgah_index_t* vgah_index_build_text(const char* gfa_text, uint32_t k, uint64_t max_furcations, uint64_t max_degree);
vgah_index_t* vgah_index_load(const char* path);   /* "*.idx": the reference's bincode file (idx_file.cpp); otherwise the .vgidx container */
int vgah_index_save_idx(const vgah_index_t*, const char* path);   /* writes a `.idx` (bincode layout of `Index`, see idx_file.cpp) */
int vgah_index_save(const vgah_index_t*, const char* path);
void vgah_index_free(vgah_index_t*);
int vgah_index_view(const vgah_index_t*, VgaIndexView* view);       /* borrowed pointers into the index */
const char* vgah_index_keys(const vgah_index_t*, uint64_t* n_keys); /* n_keys*k bytes: the k-mer strings */
const char* vgah_index_seq_rev(const vgah_index_t*);                /* Index.seq_rev */

/* Key derivation (`Index::walk_kmer`): the 2-bit packed k-mer of every key, recovered from what a `.idx`
 * holds — the key's first table position (key_offset, tab_start/tab_end), seq_fwd, node_ref, edges — plus the key's own hash
 * to tell apart k-mers that share their first and last position (the alleles of a bubble).  view->key_kmer2bit is not read.
 * hash: the function the keys were made with (Rust host: generate_hash, reference src/kmer.rs:931-934); NULL: no hash function on this
 * host — ambiguous keys are told apart by their rank among the keys that share the position pair (the table is in sorted k-mer order).
 * n_ambiguous: keys whose (start, end) admitted more than one k-mer; n_hopping: forward keys with end - start != k. */
typedef uint64_t (*vgah_hash_fn)(const char* kmer, uint32_t k, void* user);
int vgah_derive_keys(const VgaIndexView* view, const uint64_t* key_hash, vgah_hash_fn hash, void* user,
                     uint64_t* key_kmer2bit_out, uint64_t* n_ambiguous, uint64_t* n_hopping);

/* read_seqs_from_file (reference src/io.rs:74-131): extension decides FASTA/FASTQ */
vgah_reads_t* vgah_reads_from_file(const char* path);
/* seeded read simulator: forward random walks over the graph, substitutions / insertions /
 * deletions, a fraction reverse-complemented (SURVEY.md §8d inputs) */
vgah_reads_t* vgah_reads_simulate(const vgah_index_t*, uint32_t n, uint32_t len, double sub, double ins, double del,
                                  double rev_frac, uint64_t seed);
/* the same generator with one RNG stream per read (seeded by seed and first_id + i), spread over `threads` host threads
 * (0 = all cores): any sub-range of a read set can be produced independently.  Names are "read<i>". */
vgah_reads_t* vgah_reads_simulate_mt(const vgah_index_t*, uint64_t first_id, uint32_t n, uint32_t len, double sub, double ins,
                                     double del, double rev_frac, uint64_t seed, int threads);
void vgah_reads_free(vgah_reads_t*);
uint32_t vgah_reads_count(const vgah_reads_t*);
const char* vgah_reads_seqs(const vgah_reads_t*);        /* concatenated */
const uint64_t* vgah_reads_offsets(const vgah_reads_t*); /* n+1 */
const char* vgah_reads_names(const vgah_reads_t*);       /* '\n' separated */
int vgah_reads_write_fasta(const vgah_reads_t*, const char* path);

/* synthetic variation graph (config C4): random backbone, ~1 variant site per `site_every` bp,
 * 80% SNP bubbles, 10% insertions, 10% deletions; GFA text, ids topologically ordered */
char* vgah_synth_gfa(uint64_t backbone_len, uint32_t site_every, uint64_t seed, uint64_t* len);
void vgah_free_str(char*);

/* GAF text from the C ABI's result arrays: GAFAlignment::{from_chain, from_placeholder_chain,
 * to_string} (src/align.rs:762-1027) and generate_alignment (src/align.rs:1096-1168).
 * which: 0 = chains file, 1 = alignments file, 3 = alignments file with the cs string without its "cs:Z:" tag. */
char* vgah_gaf(const vgah_index_t*, const VgaResultsView* res, const vgah_reads_t* reads, int which, uint64_t* len);
/* Same from plain arrays, formatted on `threads` host threads (0 = all cores): names_joined is '\n' separated or NULL
 * ("read<name_base+i>"); *n_failed counts reads whose device alignment did not complete (status other than OK or
 * PLACEHOLDER) — they are written as placeholder records, the reference has no record type for them. */
char* vgah_gaf_raw(const vgah_index_t*, const VgaResultsView* res, const char* names_joined, uint64_t name_base, const char* seqs,
                   const uint64_t* seq_off, int which, int threads, uint64_t* len, uint64_t* n_failed);

#ifdef __cplusplus
}
#endif
#endif
