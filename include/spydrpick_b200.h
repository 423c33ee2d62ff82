/* spydrpick_b200.h -- C-ABI of the B200-native SpydrPick hot path.
 *
 * The reference (santeripuranen/SpydrPick @ f6bc110) has no plugin / FFI boundary; the
 * accelerated path sits behind four C++ template entry points called only from main():
 *
 *   apegrunt::cache_sample_weights()    externals/apegrunt/include/apegrunt/Alignment_utility.hpp:186
 *     -> apegrunt::calculate_weights()  externals/apegrunt/include/apegrunt/Alignment_StateVector_weights.hpp:192
 *   spydrpick::determine_MI_threshold() include/SpydrPick.hpp:285
 *   spydrpick::get_MI_network()         include/SpydrPick.hpp:132
 *   aracne::run_ARACNE()                include/ARACNE.hpp:550
 *
 * Each function below names the reference interface it replaces. Conventions:
 *   - plain pointers and sizes only; no C++ / torch types; no exceptions; no exit().
 *   - return 0 (SPK_OK) or a negative SPK_E_* code; message via spk_last_error().
 *   - states are the reference's codes a=0,t=1,c=2,g=3,GAP=4
 *     (externals/apegrunt/src/StateVector_state_types.cpp:33-44), dense row-major [n][L].
 *   - one host thread drives a context; calls are blocking unless stated; not re-entrant.
 *   - there is NO CPU fallback: every entry point fails with SPK_E_CUDA when no sm_100 device
 *     is usable.
 */
#ifndef SPYDRPICK_B200_H
#define SPYDRPICK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SPK_OK            0
#define SPK_E_INVALID    -1   /* bad argument / call order */
#define SPK_E_CUDA       -2   /* CUDA runtime / driver error, or no usable device */
#define SPK_E_NOMEM      -3   /* device or host allocation failed */
#define SPK_E_OVERFLOW   -4   /* candidate / edge buffers could not be grown to the size the scan needs */
#define SPK_E_PRECISION  -5   /* fp32 filter error exceeded the safety band (raise n_terms / band) */

typedef struct spk_ctx spk_ctx;

/* Edge record == apegrunt::Edge (externals/apegrunt/include/graph/Graph.hpp:113-147):
 * i < j are filtered-column indices, mi the weight. 16 bytes. */
typedef struct { uint32_t i, j; double mi; } spk_edge;

/* ---- context ------------------------------------------------------------------------ */
/* device < 0: use the current CUDA device. */
int         spk_create(spk_ctx** out, int device, int verbose);
void        spk_destroy(spk_ctx* ctx);
const char* spk_last_error(const spk_ctx* ctx);     /* ctx may be NULL: last create error */
/* Run all work of this context on an externally owned cudaStream_t (e.g. torch's current
 * stream). NULL restores the context's own stream. */
int         spk_set_stream(spk_ctx* ctx, void* cuda_stream);

/* Page-locked host memory for the caller's large buffers (the alignment): spk_set_alignment copies from
 * pinned memory at full PCIe rate; pageable memory works too (the driver stages it), only slower. */
int  spk_host_alloc(size_t bytes, void** out);
void spk_host_free(void* p);

/* ---- alignment upload ---------------------------------------------------------------- */
/* Replaces the reference's in-memory Alignment as seen by the hot path
 * (apegrunt/Alignment_interface.hpp:62-207): dense states + loci translation pos[L]
 * (original positions, apegrunt/Distance.hpp:63-83). Copies to the device and builds the
 * packed forms (bit planes, column-major bytes, state-presence masks =
 * apegrunt/Alignment_impl_block_compressed_storage.hpp:1089-1128). pos may be NULL (= 0..L-1).
 * Weights are reset to 1 (StateVector default weight). */
int spk_set_alignment(spk_ctx* ctx, const uint8_t* states, uint32_t n, uint32_t L, const uint64_t* pos);
/* Same, with the dense states already resident in the context's device memory (row-major uint8[n][L]); pos stays a
 * host array. For the one-process-per-GPU layout: every rank uploads n/G rows over its own PCIe link and the slices
 * are all-gathered over NVLink ("alignment broadcast once") instead of G full host-to-device copies.
 * The buffer must be complete when the call is made (it is read on the context's stream, which is not ordered after
 * the caller's other streams); it is copied, and the caller may free it after the call returns. */
int spk_set_alignment_device(spk_ctx* ctx, const void* d_states, uint32_t n, uint32_t L, const uint64_t* pos);

/* ---- K1: sample reweighting ----------------------------------------------------------- */
/* Column filter pre-pass. Replaces apegrunt::Alignment_filter's frequency scan + column extraction
 * (externals/apegrunt/include/apegrunt/Alignment_filter.hpp:250-339, called from src/SpydrPick.cpp:234-297).
 * spk_column_state_counts: counts[L*5] = unweighted number of samples in state a of column c (counts[c*5 + a]) of the RAW
 *   alignment states[n][L] (host, row-major, codes 0..4).
 * spk_set_alignment_filtered: uploads the raw alignment ONCE, counts on the device, applies the reference's rule on the
 *   host in its arithmetic (keep iff more than one non-gap state has frequency >= maf_threshold and the gap frequency
 *   is <= gap_threshold), extracts the kept columns on the device and makes them the context's alignment (as
 *   spk_set_alignment would with the filtered matrix and positions). keep[L] receives 1/0 per raw column, *n_kept the
 *   number kept; fewer than 2 kept columns is SPK_E_INVALID. pos may be NULL (= column index). */
int spk_column_state_counts(spk_ctx* ctx, const uint8_t* states, uint32_t n, uint64_t L, uint32_t* counts);
int spk_set_alignment_filtered(spk_ctx* ctx, const uint8_t* states, uint32_t n, uint64_t L, const uint64_t* pos,
                               double maf_threshold, double gap_threshold, uint8_t* keep, uint64_t* n_kept);

/* Replaces apegrunt::calculate_weights (Alignment_StateVector_weights.hpp:192-353) up to the
 * integer neighbour counts: counts[i] = 1 + #{j != i : similar(i,j)} with
 * similar = ident(i,j) >= min_identical, min_identical = (size_t)(threshold*(double)L) computed
 * by the caller as the reference does (:228,251). `quirks` is a mask of the reference behaviours to reproduce
 * (both pinned against the compiled reference, tests/test_ref_pins.py; SPK_QUIRKS_REFERENCE = what the reference does):
 *   SPK_QUIRK_LASTBLOCK  is_similar_to_impl counts the final 32-column block twice when L % 32 == 0
 *                        (StateVector_impl_block_compressed_alignment_storage.hpp:628-659);
 *   SPK_QUIRK_SAMPLE0    calculate_weights compares sample 0 with every sample (itself included) in its cache-priming pass
 *                        (range_primer, Alignment_StateVector_weights.hpp:218-238, 275-283) and then again in the triangular
 *                        pass (:240-262): counts[0] = 3 + 2 * #{j > 0 similar to 0}, and every j > 0 similar to sample 0
 *                        carries one extra count. */
#define SPK_QUIRK_LASTBLOCK  1
#define SPK_QUIRK_SAMPLE0    2
#define SPK_QUIRKS_REFERENCE 3
int spk_neighbour_counts(spk_ctx* ctx, uint64_t min_identical, int quirks,
                         uint64_t* counts /* [n] host */);
/* Multi-GPU sharding of K1: the sample-tile pairs are dealt round-robin to n_parts partitions; every partition
 * returns PARTIAL counts whose element-wise sum over the partitions equals spk_neighbour_counts' result
 * (partition 0 carries each sample's own multiplicity). n_parts <= 1: identical to spk_neighbour_counts. */
int spk_neighbour_counts_part(spk_ctx* ctx, uint64_t min_identical, int quirks,
                              uint32_t part_index, uint32_t n_parts, uint64_t* counts /* [n] host */);

/* K1 without any allocation. NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same result as
 * spk_set_alignment + spk_neighbour_counts (bit-exact; both quirks honoured), but context-free and on the CURRENT device:
 *   d_states  device, row-major uint8[n][L], codes 0..4          d_counts  device, uint64[n], written by the call
 *   cuda_stream  the cudaStream_t to launch on (NULL = default stream); the call only enqueues work and returns.
 * The bit planes live in a fixed file-scope __device__ array of SPK_NOALLOC_MAX_SCRATCH_BYTES that every GPU loading the library
 * holds until exit: 3 * n * ceil(L / 32) * 4 bytes must not exceed it, larger problems are rejected with SPK_E_INVALID (message
 * via spk_last_error(NULL)). The call makes no cudaMalloc / cudaFree / allocating library call. All calls on a device share the
 * array (each call overwrites the extent it uses): calls that overlap on the GPU race -- issue them on one stream or order them
 * with events. The first call on a device resolves the array's address (cudaGetSymbolAddress); make it before a graph capture. */
#define SPK_NOALLOC_MAX_SCRATCH_BYTES (16u << 20)
int spk_neighbour_counts_noalloc(const void* d_states, uint32_t n, uint32_t L, uint64_t min_identical, int quirks,
                                 uint64_t* d_counts, void* cuda_stream);

/* K1 + the computed branch of cache_sample_weights in one call (Alignment_utility.hpp:219-251): neighbour counts -> w = 1 / count
 * -> optional rescale (w * n) / n_eff (:245) -> the context's weights. weights_out / counts_out ([n], host) may be NULL.
 * On a context that belongs to a communicator this is a collective call (tile pairs sharded, partial counts all-reduced). */
int spk_reweight(spk_ctx* ctx, uint64_t min_identical, int quirks, int rescale, double* weights_out, uint64_t* counts_out);

/* Replaces the per-sequence set_weight() done by cache_sample_weights
 * (Alignment_utility.hpp:186-258) / MI_Parameters' weight snapshot (include/mi_parameters.hpp:53-59). */
int spk_set_weights(spk_ctx* ctx, const double* w /* [n] host */);

/* ---- K2b: MI for an explicit pair list ------------------------------------------------ */
/* Replaces MI_solver::single over sampled pairs (include/mi.hpp:183-224,367-372;
 * include/SpydrPick.hpp:210-255). fp64, the reference's exact formula. */
int spk_mi_pairs(spk_ctx* ctx, const uint32_t* pi, const uint32_t* pj, uint64_t npairs,
                 double pseudocount, double* mi /* [npairs] host */);

/* Replaces spydrpick::determine_MI_threshold (include/SpydrPick.hpp:285-343). The reference
 * seeds std::mt19937 from the clock (:178); here the seed is explicit (seed+iteration). */
int spk_determine_mi_threshold(spk_ctx* ctx, uint64_t n_values, uint64_t threshold_pairs /*0=auto*/,
                               uint32_t iterations, double pseudocount, uint64_t seed, double* threshold);
/* The same estimator with a choice of pair sampler. Either way all iterations go through the device together (class-bucketed fp64
 * kernels + two global radix sorts); only `iterations` doubles come back. n_values >= L(L-1)/2 is SPK_E_INVALID (the reference indexes out of range).
 *   SPK_SAMPLER_REFERENCE  the reference's sample_pairs() (std::mt19937(seed + iteration), duplicates rejected; SpydrPick.hpp:171-207),
 *                          one host thread per iteration -- what spk_determine_mi_threshold uses; reproducible against the oracle
 *   SPK_SAMPLER_DEVICE     distinct pairs from a keyed Feistel permutation of the pair index space, generated on the device */
#define SPK_SAMPLER_REFERENCE 0
#define SPK_SAMPLER_DEVICE    1
int spk_determine_mi_threshold_ex(spk_ctx* ctx, uint64_t n_values, uint64_t threshold_pairs, uint32_t iterations,
                                  double pseudocount, uint64_t seed, int sampler, double* threshold);

/* ---- K2: all-pairs scan --------------------------------------------------------------- */
typedef struct {
    double   threshold;        /* strict: edge saved iff threshold < mi          (include/mi.hpp:430,454) */
    double   pseudocount;      /* --mi-pseudocount, default 0.5                  (src/SpydrPick_options.cpp:41) */
    uint64_t ld_threshold;     /* --ld-threshold; colmax uses distance > this    (include/mi.hpp:423,447) */
    int32_t  linear_genome;    /* 0 = circular (default), 1 = linear             (include/mi.hpp:313-320) */
    uint64_t circle_size;      /* 0 = L (the reference's choice, apegrunt/Distance.hpp:67-71) */
    uint64_t max_edges;        /* capacity hint for saved edges; 0 = 2*min(1e7,100*L)+slack. A scan that needs more grows its
                                  buffers and repeats itself (the reference stores any number of edges) */
    int32_t  sort_desc;        /* 1: edges sorted (weight desc, id asc) = Graph::sort (graph/Graph.hpp:234-242) */
    int32_t  n_terms;          /* engine 2 only: bf16 terms of the weight split, 1..3, 0 = default (2) */
    double   band;             /* filter safety band delta (candidates: mi32 > threshold - band); 0 = default */
    int32_t  engine;           /* tensor-core filter + fp64 refine: 0 = default int8 engine (3 or 4);
                                  3 = int8 compact one-hot GEMM (tcgen05 kind::i8, exact s32 counts), 16-bit weights as
                                      {hi, lo} byte rows of the N operand;
                                  4 = the same with factorised weights q = a*b (a in the M operand, b in the N operand):
                                      half the tensor work per column pair, weights on the coarser byte-product grid;
                                  2 = dense 4-state bf16 GEMM with hi/lo weight split;
                                  1 = fp64 exhaustive (diagnostic, no tensor cores) */
    uint32_t part_index;       /* tile-space partition: this context scans part_index of n_parts (caller-driven sharding) */
    uint32_t n_parts;          /* 1 = whole upper triangle. 0 = automatic: the whole triangle on a single-device context; on a context
                                  that belongs to a communicator (spk_comm_*) the COLLECTIVE scan: this rank's partition, then the
                                  exchange (edge lists all-gathered and merged, colmax max-reduced, MI-without-gaps gathered on rank 0) */
    int32_t  keep_on_device;   /* 0: edges copied to pinned host buffers before the call returns;
                                  1: no host copy (device pointers below stay valid);
                                  2: host copy started on a side stream: the host pointers are returned at once and
                                     are valid after spk_wait_download(); spk_aracne() on the returned edges pointer
                                     may run in between (it uses the device copy), so the download overlaps K3 */
} spk_scan_params;

typedef struct {
    const spk_edge* edges;         uint64_t n_edges;     /* host (NULL if keep_on_device) */
    const spk_edge* edges_wo_gaps; uint64_t n_wo_gaps;   /* MI without gaps for saved edges touching a gap column
                                                            (include/mi.hpp:465-489, literal semantics) */
    const double*   colmax;                              /* [L] per-column max MI (include/mi.hpp:245-290), host */
    const void*     d_edges;                             /* device copies (spk_edge[], spk_edge[], double[L]) */
    const void*     d_edges_wo_gaps;
    const void*     d_colmax;
    uint64_t n_candidates;       /* pairs refined in fp64 */
    double   max_filter_error;   /* max |mi_fp32 - mi_fp64| over the candidates */
    /* Audit of the tensor-core filter (int8 engines): an UNBIASED sample of pairs on both sides of the candidate cut-off
     * (every pair of a 1/32 sample of the tiles with mi_fp32 > threshold - 4*band) is recomputed in fp64. A pair with
     * mi_fp64 > threshold whose mi_fp32 was not above threshold - band would have been LOST by the filter: any such pair, or an
     * audit error above band/2, makes the scan repeat with a wider band (band = 0) or fail with SPK_E_PRECISION (band given). */
    uint64_t audit_pairs;        /* pairs audited (this rank's share in a collective scan) */
    uint64_t audit_violations;   /* audited pairs the filter would have lost (0 in any result that is returned) */
    double   audit_max_error;    /* max |mi_fp32 - mi_fp64| over the audited pairs */
    double   band;               /* the band the returned result was produced with */
} spk_scan_result;

/* Replaces spydrpick::get_MI_network up to (not including) the quartile step
 * (include/SpydrPick.hpp:132-168; MI_solver::operator() include/mi.hpp:377-508).
 * Result memory is owned by the context and valid until the next scan / destroy. */
int spk_mi_scan(spk_ctx* ctx, const spk_scan_params* params, spk_scan_result* result);

/* Tukey fences from per-column maxima: maxvaltracker::quartile + get_MI_network
 * (include/mi.hpp:267-281, include/SpydrPick.hpp:148-152). Host arithmetic. */
int spk_outlier_thresholds(const double* colmax, uint64_t L, double* outlier, double* extreme);

/* ---- sort + K3: ARACNE ----------------------------------------------------------------- */
/* Graph::sort (graph/Graph.hpp:234-242): weight desc, then id = (node2<<32)|node1 asc. In place, host array. */
int spk_sort_edges(spk_ctx* ctx, spk_edge* edges, uint64_t n_edges);

/* Replaces aracne::run_ARACNE (include/ARACNE.hpp:411-550). edges must be sorted as above.
 * When `edges` is the pointer spk_mi_scan returned (sort_desc = 1, list unchanged) the device copy of the
 * scan is reused and nothing is uploaded.
 * direct_flags[e] = 1 for DIRECT edges (the reference's bool(edge), the ARACNE output column),
 * 0 for edges with a triangle in which they are the weakest and mid-min >= edge_threshold.
 * Edges [part_index*E/n_parts, (part_index+1)*E/n_parts) are evaluated (others left untouched);
 * n_parts <= 1 evaluates all. */
int spk_aracne(spk_ctx* ctx, const spk_edge* edges, uint64_t n_edges, double edge_threshold,
               uint32_t part_index, uint32_t n_parts, uint8_t* direct_flags /* [n_edges] host */);

/* Blocks until the host copies started by spk_mi_scan(keep_on_device = 2) are complete (no-op otherwise). */
int spk_wait_download(spk_ctx* ctx);

/* Device-resident variants for callers that keep data in HBM (bench, multi-GPU plumbing):
 * d_edges is a device pointer to n_edges sorted spk_edge records, d_flags device uint8[n_edges]. */
int spk_sort_edges_device(spk_ctx* ctx, void* d_edges, uint64_t n_edges);
/* d_edges holds n_runs consecutive runs of run_counts[r] edges, each already in Graph::sort order (graph/Graph.hpp:234-242;
 * e.g. the per-rank lists of a sharded scan, each sorted by its own GPU, after the all-gather): merged in place into the
 * one sorted list -- the same list spk_sort_edges_device gives, for G-1 merges instead of a full sort on every rank. */
int spk_merge_sorted_runs_device(spk_ctx* ctx, void* d_edges, const uint64_t* run_counts, uint32_t n_runs);
int spk_aracne_device(spk_ctx* ctx, const void* d_edges, uint64_t n_edges, double edge_threshold,
                      uint32_t part_index, uint32_t n_parts, void* d_flags);

/* ---- multi-GPU: one context per device, SPMD -------------------------------------------------------------------------
 * A context can belong to an NCCL communicator (one rank per context / device). The calls below that say "collective" must then
 * be made by EVERY rank's context with the same arguments -- from one host thread per context (single process, spk_comm_init_all)
 * or from one process per GPU (spk_comm_unique_id on rank 0, the 128 bytes passed to the others out of band, spk_comm_init_rank).
 * Collective calls: spk_set_alignment (rows uploaded per rank + all-gather), spk_reweight / spk_neighbour_counts,
 * spk_determine_mi_threshold[_ex] (iterations dealt to the ranks), spk_mi_scan with n_parts = 0, spk_aracne / spk_aracne_device
 * with n_parts = 0. This replaces the TBB split / join of the reference (include/SpydrPick.hpp:141-146, include/mi.hpp:338-361)
 * across devices. NCCL (libnccl.so.2) is loaded on first use. */
int spk_comm_init_all(spk_ctx** ctxs, int n);                       /* single process: n contexts on n distinct devices */
int spk_comm_unique_id(void* id128 /* 128 bytes out */);
int spk_comm_init_rank(spk_ctx* ctx, const void* id128, int rank, int world);
int spk_comm_rank(const spk_ctx* ctx, int* rank, int* world);

/* ---- instrumentation -------------------------------------------------------------------- */
/* Device time (CUDA events on the context's stream) of the stages of the most recent calls, ms. */
typedef struct {
    float upload_ms, prep_ms;               /* spk_set_alignment: H2D copy, packing kernels */
    float hamming_ms;                       /* K1 kernel */
    float expand_ms, stats_ms;              /* operand expansion, column statistics */
    float gemm_ms;                          /* K2 tcgen05 filter kernel */
    float refine_ms;                        /* K2r/K2c fp64 refine kernel */
    float compact_ms, sort_ms;              /* threshold/colmax compaction, edge sort */
    float aracne_build_ms, aracne_ms;       /* CSR build, K3 kernel */
    float pairs_ms;                         /* K2b kernel */
    uint64_t gemm_tiles, kernel_launches;   /* tiles processed by the last scan; kernels launched since create */
    uint64_t gemm_pairs;                    /* column pairs covered by those tiles (incl. padding / masked diagonal halves) */
    float    seed_ms;                       /* int8 engine: the colmax seed launch, included in gemm_ms (main pass = gemm_ms - seed_ms) */
    float    exchange_ms;                   /* collective scan: counts + edge all-gather + colmax all-reduce + run merge */
    float    audit_ms;                      /* filter audit: appending the audit sample to the candidate list (its fp64 recomputation is part of refine_ms) */
    float    threshold_ms;                  /* threshold estimator: sampling + K2b + sorts (device part) */
} spk_timings;
int spk_get_timings(spk_ctx* ctx, spk_timings* out);

/* Debug/validation hook: raw fp32 joint counts of one 32x64 column tile as produced by the
 * tcgen05 mainloop (counts[32][64][4][4], index [i][j][a][b]); used by the parity tests to
 * localise descriptor / layout faults. */
int spk_debug_tile_counts(spk_ctx* ctx, uint32_t iblock32, uint32_t jblock64, int32_t n_terms, float* counts);

/* int8 compact engine introspection (tests): tiles per class {1,2,4 rows}, first permuted column of each
 * class, the permutation (original column per permuted column, 0xffffffff = padding), per-column info
 * (k | kept state codes, 3 bits each from bit 4 | dropped state << 16), quantised 16-bit weights and the
 * weight per quantisation step. perm / info / q / delta may be NULL. */
/* weight rows of the default int8 engine: 2 = engine 3 ({hi, lo} bytes), 1 = engine 4 (factorised weights) */
int spk_debug_i8_mode(void);
/* Measured peak of the tensor instruction the int8 filter issues (tcgen05.mma.cta_group::2.kind::i8 M256 N256 K32 on operands already in
 * shared memory, one CTA pair per TPC, no loads, no epilogue): `reps` launches of `kblocks` 128-byte K blocks each, back to back.
 * *tops = u8 tensor ops per second (1e12) over the interval, *ms = its device time. bench.py's roofline denominator. */
int spk_debug_i8_peak(spk_ctx* ctx, uint32_t kblocks, uint32_t reps, double* tops, float* ms);
int spk_debug_i8_layout(spk_ctx* ctx, uint32_t* ntiles3, uint32_t* col_start3, uint32_t* perm, uint32_t* info,
                        uint64_t perm_cap, uint32_t* q /* [n] */, double* delta);
/* Raw s32 joint counts (quantised-weight units) of the tile pair (ti <= tj) as produced by the kind::i8
 * mainloop: tables[(128/KI)][(128/KJ)][KI][KJ], KI/KJ = rows per column of the two tiles' classes. */
int spk_debug_i8_tile(spk_ctx* ctx, uint32_t ti, uint32_t tj, uint32_t* tables /* [16384] */);

const char* spk_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SPYDRPICK_B200_H */
