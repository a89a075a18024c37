/*
 * seqrush_b200.h -- C ABI of the B200-native align-and-unite hot path of pangenome/seqrush.
 *
 * Plain C, plain pointers and sizes; this is what a Rust `extern "C"` block (or cgo / ctypes)
 * binds. Each entry point cites the reference interface it replaces (paths relative to the
 * seqrush repository). All functions return 0 on success and a negative SRGPU_E_* code on
 * failure unless stated otherwise; srgpu_last_error() gives a message for the calling thread.
 *
 * Threading: every entry point may be called concurrently from several host threads (the
 * reference calls BidirectedUnionFind through `&self` from rayon workers, src/seqrush.rs:738-756);
 * calls on one context are serialised internally.
 */
#ifndef SEQRUSH_B200_H_
#define SEQRUSH_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRGPU_OK 0
#define SRGPU_E_CUDA (-1)      /* CUDA runtime error (no device, OOM, launch failure) */
#define SRGPU_E_ARG (-2)       /* invalid argument (index out of range, bad penalties) */
#define SRGPU_E_RANGE (-3)     /* Pos outside the union-find (reference: assert in uf_rush lib.rs:113) */
/* (-4 was SRGPU_E_VALIDATE, the reference's validate_match panic, src/seqrush.rs:1184-1205. No entry point can
 * produce it: every CIGAR walk builds its match runs from bases that compared equal, so it was removed.) */
#define SRGPU_E_INTERNAL (-5)

/* Per-pair status in srgpu_result.status */
#define SRGPU_PAIR_OK 0
#define SRGPU_PAIR_FAILED 1 /* reference: alignment status != Completed -> empty CIGAR, score i32::MAX (allwave/src/alignment.rs:40-55) */

typedef struct srgpu_ctx srgpu_ctx;
typedef struct srgpu_uf srgpu_uf;

/* Borrowed view of one sequence (reference: allwave::Sequence{id,seq}, allwave/src/types.rs:6-10;
 * seqrush Sequence{id,data,offset}, src/seqrush.rs:272-277). Offsets in the concatenated Pos
 * space are the running sum of lengths in array order (src/seqrush.rs:1816-1818). */
typedef struct {
  const uint8_t* data;
  uint64_t len;
} srgpu_seqview;

/* Reference: allwave::AlignmentParams (allwave/src/types.rs:37-46) in any of allwave's modes
 * (AlignmentMode::from_params, types.rs:101-111; allwave/src/alignment.rs:203-231):
 *   two-piece affine    gap_ext2 > 0;
 *   single-piece affine gap_open2 / gap_ext2 <= 0 (the Option is None): the gap-affine BiWFA;
 *   "edit distance"     the same with gap_open1 == gap_ext1 == mismatch (allwave runs it as affine x,x,x).
 * match_score follows WFA2-lib: 0, or negative with Eizenga's adjustment of the other penalties
 * (WFA2-lib/wavefront/wavefront_penalties.c:113-134,157-175; the score reported is then
 * (|match| * (qlen + tlen) - wf_score) / 2, wavefront_compute.c:108-120). A positive match score is refused with
 * SRGPU_E_ARG and WFA2-lib's message: WFA2-lib itself prints it and exits (:100-103,141-144). */
typedef struct {
  int32_t match_score;
  int32_t mismatch;
  int32_t gap_open1;
  int32_t gap_ext1;
  int32_t gap_open2;
  int32_t gap_ext2;
} srgpu_penalties;

/* Reference: allwave::AlignmentResult (allwave/src/types.rs:14-35). CIGAR bytes are in WFA
 * convention ('M','X','I' = consumes target, 'D' = consumes query), one byte per column, so the
 * reference's cigar_bytes_to_string (allwave/src/alignment.rs:327-356) applies unchanged. */
typedef struct {
  uint32_t query_idx;
  uint32_t target_idx;
  int32_t is_reverse;
  int32_t status;            /* SRGPU_PAIR_* */
  int32_t score;             /* wf.score(): -penalty when BiWFA split at level 0, INT32_MIN when the
                                whole pair went through the base case (WFA2-lib quirk), INT32_MAX on failure */
  int32_t penalty;           /* gap-affine-2p penalty recomputed from the CIGAR (always meaningful) */
  uint32_t num_matches;      /* count of 'M' */
  uint32_t alignment_length; /* count of 'M' + 'X' (allwave/src/alignment.rs:279-301) */
  uint32_t query_end;        /* M + X + D */
  uint32_t target_end;       /* M + X + I */
  uint32_t edit_distance;    /* min(fwd, rc) unit-cost distance found by the orientation step */
  uint32_t n_runs;           /* run-length CIGAR elements */
  uint64_t cigar_off;        /* offset into the arena returned by srgpu_align_pairs */
  uint64_t cigar_len;
  uint64_t wf_cells;         /* wavefront cells computed for this pair (5-component cells) */
} srgpu_result;

/* Counters and device timings of the last run on a context. */
typedef struct {
  uint64_t pairs;
  uint64_t aligned_bp;     /* sum of query lengths */
  uint64_t wf_cells;       /* BiWFA cells (one diagonal at one score, 5 components) */
  uint64_t orient_cells;   /* unit-cost cells of the orientation step */
  uint64_t unite_ops;      /* BidirectedUnionFind::unite calls issued by unite_matching_region */
  uint64_t unite_links;    /* unions that actually linked two roots */
  uint64_t runs;           /* run-length CIGAR elements produced */
  uint32_t kernel_launches;
  uint32_t align_launches;
  uint32_t union_launches;
  uint32_t retries;
  float align_ms;          /* CUDA-event time of the align kernel launches */
  float union_ms;          /* CUDA-event time of the union kernel launches */
  float total_ms;
  float orient_ms;         /* CUDA-event time of the strand-choice (orient) kernel launches */
  uint32_t failed_pairs;   /* pairs the reference would fail too (WF_STATUS_END_UNREACHABLE -> empty CIGAR, score
                              i32::MAX, allwave/src/alignment.rs:40-55); they are skipped by the union step. Pairs
                              that exceed a workspace limit of THIS library never end up here: the call fails with
                              SRGPU_E_INTERNAL instead of returning a partition that silently lacks them */
  uint32_t reserved0;
} srgpu_stats;

const char* srgpu_last_error(void);

/* Context: one per GPU. `stream` is a cudaStream_t (NULL = the library creates its own). */
int srgpu_ctx_create(int device, void* stream, srgpu_ctx** out);
void srgpu_ctx_destroy(srgpu_ctx* ctx);
int srgpu_ctx_synchronize(srgpu_ctx* ctx);
/* The cudaStream_t every launch of this context goes to (so a host can order its own work, e.g. an
 * NCCL collective, against it without a host synchronisation). */
void* srgpu_ctx_stream(srgpu_ctx* ctx);
/* Tunables: threads per CTA (multiple of 32), CTAs per SM, initial ring width in diagonals
 * (0 = automatic). */
int srgpu_ctx_configure(srgpu_ctx* ctx, int threads, int ctas_per_sm, int ring_diagonals);
/* Tunable of the alignment launches (no reference counterpart): BiWFA sub-problems whose remaining score is at
 * most `split_score` are finished by one-warp CTAs in a second launch and joined by a third; 0 keeps every
 * pair in one CTA from start to end ("monolithic"), values below 250 (the reference's base-case threshold,
 * WFA2-lib/wavefront/wavefront_bialign.c) are refused. Results do not depend on it. Default 1000. */
int srgpu_ctx_set_split(srgpu_ctx* ctx, int split_score);
int srgpu_ctx_stats(srgpu_ctx* ctx, srgpu_stats* out);

/* Upload the sequence set (host -> HBM). Replaces the `&[Sequence]` argument of
 * AllPairIterator::with_options (allwave/src/iterator.rs:30) and SeqRush.sequences. */
int srgpu_set_sequences(srgpu_ctx* ctx, const srgpu_seqview* seqs, uint32_t n_seqs);
uint64_t srgpu_total_length(srgpu_ctx* ctx);

/* Batched replacement of AllPairIterator(...).into_par_iter() -> AlignmentResult
 * (allwave/src/iterator.rs:30-124,185-203 + allwave/src/alignment.rs:16-57): orientation by exact
 * edit distance, then BiWFA gap-affine-2p, end-to-end, no heuristic. pairs[i] = {query, target}.
 * On return *cigar_arena (release with srgpu_free) holds all CIGAR bytes; out[i] indexes into it.
 * cigar_arena may be NULL to skip CIGAR expansion (results only). */
int srgpu_align_pairs(srgpu_ctx* ctx, const uint32_t (*pairs)[2], uint64_t n_pairs, const srgpu_penalties* pen,
                      srgpu_result* out, uint8_t** cigar_arena, uint64_t* cigar_arena_len);
void srgpu_free(void* p);

/* Fused replacement of the closure body src/seqrush.rs:738-756 (align_pair + cigar_bytes_to_string
 * + process_alignment): align every pair, walk its CIGAR on the device and unite the exact-match
 * runs of length >= min_match_len into `uf`. out may be NULL. */
int srgpu_align_and_unite(srgpu_ctx* ctx, srgpu_uf* uf, const uint32_t (*pairs)[2], uint64_t n_pairs,
                          const srgpu_penalties* pen, uint64_t min_match_len, srgpu_result* out);
/* Same with the pair list already resident in HBM (device pointer to n_pairs x 2 uint32). */
int srgpu_align_and_unite_device(srgpu_ctx* ctx, srgpu_uf* uf, const void* d_pairs, uint64_t n_pairs,
                                 const srgpu_penalties* pen, uint64_t min_match_len);

/* srgpu_align_pairs that also unites: ONE alignment pass yields the CIGARs (for the PAF of
 * --output-alignments) and feeds the union-find, where the reference aligns every pair twice
 * (src/seqrush.rs:693-716 then :728-756). The unions run on the device from the run-length CIGARs
 * before they are expanded to bytes. */
int srgpu_align_pairs_and_unite(srgpu_ctx* ctx, srgpu_uf* uf, const uint32_t (*pairs)[2], uint64_t n_pairs,
                                const srgpu_penalties* pen, uint64_t min_match_len, srgpu_result* out,
                                uint8_t** cigar_arena, uint64_t* cigar_arena_len);

/* Replacement of process_alignment alone (src/seqrush.rs:1134-1481) for CIGARs that come from
 * elsewhere (the --paf path, src/seqrush.rs:510-609): cigar is the RLE string ("12=1X3I..."),
 * 'M' and '=' are re-checked base by base like the reference (so the reference's validate_match
 * panic, :1179-1207, has no counterpart: a run is built only from bases that compared equal).
 * The batched form takes every record of a PAF file in one call: the strings are parsed into run
 * lists on the host and walked by one multi-CTA launch. */
typedef struct {
  uint32_t query_idx, target_idx;
  int32_t query_is_rc; /* PAF strand == '-' (src/seqrush.rs:580) */
  int32_t reserved;
  uint64_t query_start, target_start; /* PAF columns 3 and 8 */
  const char* cigar;                  /* NUL-terminated RLE string of the cg:Z: tag ("" when absent, :552-555) */
} srgpu_paf_record;
int srgpu_process_alignments(srgpu_ctx* ctx, srgpu_uf* uf, const srgpu_paf_record* records, uint64_t n_records,
                             uint64_t min_match_len);
int srgpu_process_alignment(srgpu_ctx* ctx, srgpu_uf* uf, const char* cigar, uint32_t query_idx, uint32_t target_idx,
                            uint64_t min_match_len, int query_is_rc, uint64_t query_start, uint64_t target_start);

/* ---- BidirectedUnionFind (src/bidirected_union_find.rs:7-135) over Pos = offset*2 + is_rev
 * (src/pos.rs:6-87). HBM-resident parent array, lock-free atomicCAS unions. find() returns the
 * canonical representative = the minimum Pos of the component (the reference's representative is
 * schedule-dependent, uf_rush lib.rs:178; only the partition is contractual). ---- */
int srgpu_uf_new(srgpu_ctx* ctx, uint64_t max_offset, srgpu_uf** out); /* ::new, :16-24: capacity 2*max_offset+2 */
void srgpu_uf_free(srgpu_uf* uf);
uint64_t srgpu_uf_capacity(srgpu_uf* uf);
int srgpu_uf_clear(srgpu_uf* uf); /* UFRush::clear, uf_rush-0.2.1/src/lib.rs:211-216: all singletons again */
int srgpu_uf_find(srgpu_uf* uf, uint64_t pos, uint64_t* rep);             /* ::find, :27-31 */
int srgpu_uf_same(srgpu_uf* uf, uint64_t a, uint64_t b, int* same);       /* ::same, :46-54 */
int srgpu_uf_unite(srgpu_uf* uf, uint64_t a, uint64_t b);                 /* ::unite, :35-43 */
int srgpu_uf_unite_matching_region(srgpu_uf* uf, uint64_t seq1_offset, uint64_t seq2_offset, uint64_t seq1_local_start,
                                   uint64_t seq2_local_start, uint64_t match_length, int seq1_is_rc,
                                   uint64_t seq1_len); /* :60-98 */
int srgpu_uf_unite_matching_region_seq2_rc(srgpu_uf* uf, uint64_t seq1_offset, uint64_t seq2_offset,
                                           uint64_t seq1_local_start, uint64_t seq2_local_start, uint64_t match_length,
                                           int seq2_is_rc, uint64_t seq2_len); /* :103-129 */
typedef struct {
  uint64_t seq1_offset, seq2_offset, seq1_local_start, seq2_local_start, match_length, seq1_len;
  int32_t seq1_is_rc;
  int32_t seq2_rc_form; /* 0: unite_matching_region semantics, 1: _seq2_rc semantics (seq1_len = seq2_len) */
} srgpu_region;
int srgpu_uf_unite_regions(srgpu_uf* uf, const srgpu_region* regions, uint64_t n);
/* SeqRush::new's loop `unite(make_pos(i,false), make_pos(i,true))` for i in 0..total_length
 * (src/seqrush.rs:324-328). */
int srgpu_uf_unite_fr_all(srgpu_uf* uf, uint64_t total_length);
/* Bulk find: out[p] = find(p) for p in [first, first+n) (the 2-finds-per-base consumer,
 * src/bidirected_builder.rs:48-49). */
int srgpu_uf_roots(srgpu_uf* uf, uint64_t first, uint64_t n, uint64_t* out);
/* count_components (src/seqrush.rs:341-353): distinct representatives over forward Pos. */
int srgpu_uf_count_components(srgpu_uf* uf, uint64_t total_length, uint64_t* count);
/* count_components WITHOUT any allocation. This entry point has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it). Same result as srgpu_uf_count_components; the kernel's scratch (a 64-bit
 * counter and a bitmap with one bit per Pos) is a file-scope __device__ array of 4 MiB + 256 B that every GPU loading
 * the library holds until exit. Limit: union-finds of at most srgpu_uf_count_components_static_max_capacity() =
 * 33 554 432 Pos (16 777 215 bases); larger ones are refused with SRGPU_E_ARG. It calls no cudaMalloc / cudaFree or
 * allocating library routine. Every call shares the one array (it is cleared at the start of each call): calls that
 * overlap on the same GPU race -- serialise them. */
int srgpu_uf_count_components_static(srgpu_uf* uf, uint64_t total_length, uint64_t* count);
uint64_t srgpu_uf_count_components_static_max_capacity(void);
/* 64-bit order-independent checksum of the canonical partition (sum over p of mix(p, find(p))). */
int srgpu_uf_checksum(srgpu_uf* uf, uint64_t* checksum);
/* Multi-GPU merge (one process per GPU): device pointer + element count of the uint32 parent
 * array, a compress pass that makes parent[p] == find(p), and a merge pass that unites
 * (p, roots[p]) for a peer's gathered roots array (device pointer). */
void* srgpu_uf_device_ptr(srgpu_uf* uf);
int srgpu_uf_compress(srgpu_uf* uf);
int srgpu_uf_merge_roots(srgpu_uf* uf, const void* d_peer_roots, uint64_t n);
/* Stream-ordered forms for the end-of-step exchange (no host synchronisation): compress, then -- after
 * an all_gather of every rank's compressed parent array into one [world][n] device buffer -- ONE pass
 * that unites (p, gathered[r][p]) for every peer r != my_rank. */
int srgpu_uf_compress_async(srgpu_uf* uf);
int srgpu_uf_merge_gathered(srgpu_uf* uf, const void* d_gathered, uint64_t n, int world, int my_rank);

/* ---- Pair-list sparsification (allwave::SparsificationStrategy, allwave/src/types.rs:64-80; flag parsing
 * src/seqrush.rs:356-431). The reference derives sparsified pair lists from Rust's DefaultHasher (SipHash-1-3 with
 * zero keys): MinHash sketches of canonical k-mer hashes (allwave/src/mash.rs:78-107: `[u8]` framing = u64 LE length,
 * then the bytes), Mash distances, k-nearest / k-farthest neighbours (allwave/src/knn_graph.rs:10-142) and a hashed
 * filter on the String "{id_i}:{id_j}" (`str` framing = bytes, then 0xff; allwave/src/iterator.rs:256-284,
 * knn_graph.rs:145-172). The hashing, the sketch selection and the n^2 Jaccard counts run on the device; the host
 * turns the integer counts into distances and neighbour lists (seqrush_b200/sparsify.py). ---- */
uint64_t srgpu_rust_hash_bytes(const uint8_t* data, uint64_t n); /* DefaultHasher::new(); data.hash(); finish() for &[u8] */
uint64_t srgpu_rust_hash_str(const char* s);                     /* the same for a String / &str */
/* Sketches of the uploaded sequences (sketch_size smallest canonical k-mer hashes each, ascending, duplicates kept:
 * sketches_out[n][sketch_size], counts_out[n]) and, for every sequence pair, the sizes of the intersection and of the
 * union of the two sketches as sets (inter_out[n][n], uni_out[n][n], symmetric, diagonal 0). Outputs may be NULL. */
int srgpu_mash_sketch_jaccard(srgpu_ctx* ctx, uint32_t kmer_size, uint32_t sketch_size, uint64_t* sketches_out,
                              uint32_t* counts_out, uint32_t* inter_out, uint32_t* uni_out);
/* keep_out[i*n+j] = 1 when hash("{ids[i]}:{ids[j]}") / 2^64 < fraction (and i != j when exclude_self);
 * hashes_out (optional) receives the n*n hash values. */
int srgpu_pair_hash_filter(srgpu_ctx* ctx, const char* const* ids, uint32_t n, double fraction, int exclude_self,
                           uint8_t* keep_out, uint64_t* hashes_out);

/* ---- Graph build from the partition (the consumer of the union-find): the reference's
 * write_bidirected_gfa (src/bidirected_gfa_writer.rs:8-167) with --skip-sgd --skip-groom --skip-topo =
 * build_bidirected_graph_with_options (src/bidirected_builder.rs:17-289) -> compact + renumber unless
 * --no-compact (src/bidirected_ops.rs:75-478) -> validate_paths_match_sequences (src/seqrush.rs:1741-1798) ->
 * write_gfa (src/bidirected_ops.rs:880-925).
 *
 * srgpu_graph_build runs on the device: bulk canonical roots (two finds per base, :48-49), first-seen node
 * numbering by a prefix sum, node bases, oriented path steps, and the complement-aware edge set in the form of
 * each edge's first occurrence (add_edge, src/bidirected_ops.rs:813-825). The union-find must hold the
 * invariant of SeqRush::new (every forward Pos united with its reverse, src/seqrush.rs:324-328).
 *
 * Node base. The reference reads it at the offset of the union-find ROOT (:176-182), which for uf_rush depends
 * on union order and thread interleaving; for a component that mixes forward and reverse-complement members
 * that choice fixes the node's orientation. `representatives` == NULL uses this library's canonical
 * representative (the first-seen member, so the first step through a node is always '+'); otherwise
 * representatives[i] (i < total_length) is a Pos inside the component of position i whose offset supplies the
 * base -- pass the roots of a sequential uf_rush run to reproduce the reference at -t 1 byte for byte.
 *
 * L lines are written sorted by (from, to) handle; the reference writes them in HashSet iteration order, which
 * differs from process to process. */
typedef struct srgpu_graph srgpu_graph;
int srgpu_graph_build(srgpu_ctx* ctx, srgpu_uf* uf, const uint64_t* representatives, srgpu_graph** out);
/* The same object from caller-supplied parts (host only, no device): single-base nodes 1..n_nodes, per-path
 * handle lists (concatenated), stored edges as (from, to) handle pairs, and the sequences the paths must spell. */
int srgpu_graph_from_parts(uint64_t n_nodes, const uint8_t* node_bases, uint32_t n_paths, const uint64_t* path_lens,
                           const uint64_t* steps, uint64_t n_edges, const uint64_t* from_to, const srgpu_seqview* seqs,
                           srgpu_graph** out);
int srgpu_graph_compact(srgpu_graph* g); /* BidirectedGraph::compact + renumber_nodes_sequentially */
uint64_t srgpu_graph_node_count(srgpu_graph* g);
uint64_t srgpu_graph_edge_count(srgpu_graph* g);
uint64_t srgpu_graph_path_count(srgpu_graph* g);
uint64_t srgpu_graph_path_length(srgpu_graph* g, uint32_t path);
/* handles = (node_id << 1) | is_reverse (src/bidirected_graph.rs:6-48) */
int srgpu_graph_path_steps(srgpu_graph* g, uint32_t path, uint64_t* out, uint64_t cap);
int64_t srgpu_graph_node_sequence(srgpu_graph* g, uint64_t node_id, char* buf, uint64_t cap); /* length, -1 if absent */
int srgpu_graph_edges(srgpu_graph* g, uint64_t* from_to, uint64_t cap_edges); /* sorted (from, to) handle pairs */
/* validate + write; names[i] = id of sequence i (P line names). *text is released with srgpu_free. */
int srgpu_graph_gfa_text(srgpu_graph* g, const char* const* names, uint32_t n_names, char** text, uint64_t* len);
int srgpu_graph_write_gfa(srgpu_graph* g, const char* const* names, uint32_t n_names, const char* path);
int srgpu_graph_stats(srgpu_graph* g, uint64_t* uncompacted_nodes, uint32_t* rounds, uint32_t* merged_chains);
void srgpu_graph_free(srgpu_graph* g);

#ifdef __cplusplus
}
#endif
#endif /* SEQRUSH_B200_H_ */
