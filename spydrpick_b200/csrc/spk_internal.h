// spk_internal.h -- shared declarations of the CUDA implementation behind include/spydrpick_b200.h
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include "../../include/spydrpick_b200.h"

namespace spk {

constexpr int NSTATES = 5;
constexpr int GAP = 4;

// stage timing events (Ctx::ev)
enum {
    EV_UP0 = 0, EV_UP1, EV_PREP1,            // set_alignment: upload, packing
    EV_K1_0, EV_K1_1,                        // K1
    EV_EXP0, EV_EXP1, EV_STAT1,              // int8 operand expansion, column records
    EV_GEMM0, EV_SEED1, EV_GEMM1,            // K2 filter: start, end of the seed pass, end
    EV_REF0, EV_REF1, EV_CMP1,               // fp64 refine, compaction
    EV_SORT0, EV_SORT1,                      // edge sort
    EV_XCH0, EV_XCH1,                        // multi-GPU exchange + run merge
    EV_AR0, EV_AR1, EV_AR2,                  // ARACNE: CSR build, K3
    EV_PAIRS0, EV_PAIRS1,                    // K2b
    EV_AUD0, EV_AUD1,                        // filter audit
    EV_THR0, EV_THR1,                        // threshold estimator (device part)
    EV_TMP0, EV_TMP1,                        // short-lived local use (debug paths)
    EV_COUNT
};
enum {                                       // Ctx::tm_pending bits
    TMP_UPLOAD = 1u << 0, TMP_PREP = 1u << 1, TMP_K1 = 1u << 2, TMP_EXPAND = 1u << 3, TMP_STATS = 1u << 4, TMP_GEMM = 1u << 5,
    TMP_SEED = 1u << 6, TMP_REFINE = 1u << 7, TMP_COMPACT = 1u << 8, TMP_SORT = 1u << 9, TMP_XCH = 1u << 10, TMP_ARB = 1u << 11,
    TMP_AR = 1u << 12, TMP_PAIRS = 1u << 13, TMP_AUDIT = 1u << 14, TMP_THR = 1u << 15
};

// GEMM tile geometry (K2): A tile = 32 i-columns x 4 states = 128 rows (weighted, T terms),
// B tile = 64 j-columns x 4 states = 256 rows (unweighted one-hot), K block = 64 samples.
constexpr int TILE_I = 32;
constexpr int TILE_J = 64;
constexpr int BLOCK_K = 64;
constexpr int SUPER_I = 16;          // i-tiles per super-row (L2 panel reuse)

// per-column statistics consumed by the fp32 epilogue (64 bytes)
struct ColStat {
    float m[5];        // weighted marginal counts of states a,t,c,g,gap
    float h[5];        // h[k-1] = sum_{a in P} xlogx((m[a]+pc*k)/(W+pc*|P|*k)), k = |P_other| in 1..5 (natural log)
    uint32_t mask;     // 5-bit presence mask
    uint32_t npres;    // popcount(mask)
    uint32_t pos_lo, pos_hi;   // original position (loci translation)
    uint32_t pad[2];
};
static_assert(sizeof(ColStat) == 64, "ColStat must be 64 bytes");

// candidate emitted by the filter: pair + fp32 MI
struct Candidate { uint32_t i, j; float mi32; uint32_t pad; };

// ---- int8 compact engine (spk_mi_i8.cu) -----------------------------------------------------
// Columns are classed by k = (#present states - 1) one-hot rows they need (the highest-coded present
// state -- the gap when there is one -- is recovered from the marginals): class 0: k<=1 (1 row),
// class 1: k==2 (2 rows), class 2: k in {3,4} (4 rows). Columns are permuted so that classes are
// contiguous and every class segment is padded to whole 128-row tiles.
constexpr uint32_t I8_INVALID = 0xffffffffu;
struct ColRec8 {          // 80 bytes, per permuted column
    uint32_t mq[4];       // integer (quantised-weight) marginals of the kept states, kept order; 0 beyond k
    uint32_t mqd;         // integer marginal of the dropped (recovered) state
    uint32_t smq;         // sum of mq[]
    uint32_t k;           // kept rows actually used (0..4); #present states = k+1
    uint32_t orig;        // original (filtered-alignment) column index, I8_INVALID for padding
    uint32_t pos_lo, pos_hi;
    float h[5];           // h[m-1]    = sum_{a present} xlogx((mq_a + pcq*m)/(Wq + pcq*(k+1)*m)), natural log
    float invz[5];        // invz[m-1] = 1/(Wq + pcq*(k+1)*m)        (m = #present states of the partner column)
};
static_assert(sizeof(ColRec8) == 80, "ColRec8 must be 80 bytes");
struct I8Layout {
    uint32_t ncols[3] = {0, 0, 0}, ncols_pad[3] = {0, 0, 0}, ntiles[3] = {0, 0, 0};
    uint32_t col_start[3] = {0, 0, 0}, tile_start[3] = {0, 0, 0};
    uint32_t NT = 0, Lperm = 0;
};

struct Ctx {
    int device = 0;
    int verbose = 0;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    std::string err;
    int sm_count = 148;

    // alignment
    uint32_t n = 0, L = 0;
    uint32_t n_pad = 0;      // samples padded to BLOCK_K
    uint32_t L_pad = 0;      // columns padded to TILE_J
    uint32_t W32 = 0;        // 32-bit words per bit plane row = ceil(L/32)
    uint8_t*  d_states = nullptr;     // [n][L] raw
    uint8_t*  d_colmajor = nullptr;   // [L_pad][n_pad] column-major bytes (pad samples = 255, pad columns = GAP)
    uint32_t* d_planes = nullptr;     // [3][n][W32] bit planes
    uint64_t* d_pos = nullptr;        // [L_pad]
    uint8_t*  d_masks = nullptr;      // [L_pad]
    double*   d_w = nullptr;          // [n_pad] (pad = 0)
    double*   d_marg = nullptr;       // [L_pad][5] fp64 weighted marginals
    ColStat*  d_colstat = nullptr;    // [L_pad]
    double    w_total = 0.0;
    bool      stats_valid = false;    // d_marg / d_colstat match current weights + pseudocount
    double    stats_pc = -1.0;
    // one-hot operands
    __nv_bfloat16* d_U = nullptr;     // [4*L_pad][n_pad] unweighted
    __nv_bfloat16* d_V = nullptr;     // [T][4*L_pad][n_pad] weighted terms
    int       operands_terms = 0;     // number of valid terms in d_V (0 = invalid)
    size_t    operands_cap_terms = 0;

    // int8 compact engine
    I8Layout  i8;
    std::vector<uint8_t> h_masks;         // host copy of d_masks (layout construction)
    std::vector<double>  h_w;             // host copy of the weights (quantisation)
    uint32_t* d_i8_perm = nullptr;        // [Lperm] original column per permuted column (I8_INVALID = padding)
    uint32_t* d_i8_info = nullptr;        // [Lperm] k | kept state codes (3 bits each from bit 4) | dropped state << 16
    ColRec8*  d_i8_rec = nullptr;         // [Lperm]
    uint8_t*  d_A8 = nullptr;             // [128*NT][n_pad] one-hot u8 (M operand)
    uint8_t*  d_B8 = nullptr;             // [256*NT][n_pad] one-hot * weight byte (hi, lo) u8 (N operand)
    uint32_t* d_q = nullptr;              // [n_pad] 16-bit quantised weights
    uint32_t* d_i8_colmax32 = nullptr;    // [Lperm] running fp32 maxima, permuted index (ordered-uint)
    double    i8_delta = 0.0;             // weight represented by one quantisation step
    unsigned long long i8_wq_total = 0;   // sum of quantised weights
    bool      i8_weights_valid = false, i8_operands_valid = false;
    double    i8_rec_pc = -1.0;           // pseudocount the ColRec8 tables were built for (<0: invalid)
    uint32_t  i8_cap_Lperm = 0, i8_cap_NT = 0, i8_cap_npad = 0;   // capacities of the buffers above (grow-only per shape)

    // scan outputs (device)
    Candidate* d_cand = nullptr;  uint64_t cand_cap = 0;
    Candidate* d_cand2 = nullptr; uint64_t cand2_cap = 0;     // class-sorted copy for the fp64 refine
    spk_edge*  d_edges = nullptr; uint64_t edges_cap = 0;
    spk_edge*  d_wo = nullptr;    uint64_t wo_cap = 0;
    spk_edge*  d_refined = nullptr; double* d_refined_wo = nullptr; uint64_t refined_cap = 0;
    double*    d_colmax = nullptr;        // [L_pad] as ordered uint64 during the scan, doubles after
    float*     d_colmax32 = nullptr;      // [L_pad] running fp32 maxima (ordered-int encoded)
    unsigned long long* d_counters = nullptr;   // [8] n_cand, n_edges, n_wo, max_err_bits, ...
    // host mirrors
    spk_edge* hp_edges = nullptr; uint64_t hp_edges_cap = 0;    // pinned (cudaHostAlloc) result mirrors, grow-only
    spk_edge* hp_wo = nullptr;    uint64_t hp_wo_cap = 0;
    uint64_t  host_edges_n = 0;        // # edges of the last scan mirrored in hp_edges while d_edges still holds them (0 = stale)
    std::vector<double> h_colmax;

    // scratch
    void* d_scratch = nullptr; size_t scratch_bytes = 0;
    void* d_tilemap = nullptr;            // CUtensorMap storage is passed by value; super-row prefix table lives here
    uint64_t* d_counts = nullptr;         // [n] neighbour counts

    // weight-grouped bit planes for the fp64 kernels (spk_mi_exact.cu): samples are ordered by distinct weight
    // value, every group padded to whole 32-bit words, so N_ab = sum_w gval[w] * popc(Bi_a[w] & Bj_b[w]).
    uint32_t* d_gbits = nullptr;          // [L_pad][5][g_NW] one bit plane per (column, state)
    double*   d_gval = nullptr; uint32_t g_nw_cap = 0;           // [g_NW] weight of the samples of word w (0 for padding words)
    double*   d_gmarg = nullptr;          // [L_pad][5] fp64 weighted marginals summed the same way
    int32_t*  d_gslot = nullptr;          // [g_NW*32] sample index per bit slot, -1 = padding
    unsigned int* h_mail = nullptr; unsigned int* d_mail = nullptr;   // 64-byte page-locked mailbox, mapped into the device: small read-backs that must not queue on the D2H copy engine
    uint64_t d_edges_sorted_n = 0;        // > 0: d_edges holds this many edges of the last scan in Graph::sort order (node ids < L)
    std::vector<int32_t> h_gslot; std::vector<double> h_gval;       // host sources of the two uploads (kept alive: no sync after the async copies)
    double*   d_gH = nullptr;             // [L_pad][5] marginal entropy sum_a p_a ln p_a for partner #states 1..5
    double    g_pc = -1.0;                // pseudocount d_gH was built for
    uint32_t  g_NW = 0;                   // words per plane (multiple of 32)
    size_t    g_cap_words = 0;            // capacity of d_gbits in words
    int       g_state = 0;                // 0 = not built, 1 = built, 2 = weights not groupable (byte kernel is used)

    spk_timings tm = {};
    // Stage timing events. Every stage owns its own pair so that no host synchronisation is needed between stages: the
    // events are recorded as the work is queued, `tm_pending` remembers which intervals are outstanding and
    // resolve_timings() turns them into spk_timings fields at the next point where the stream is known to be idle.
    cudaEvent_t ev[EV_COUNT] = {};
    uint32_t tm_pending = 0;
    // multi-GPU (spk_comm.cu): NCCL communicator of this context's rank; world == 1: single device
    void* comm = nullptr; int rank = 0, world = 1;
    spk_edge* d_merge = nullptr; uint64_t merge_cap = 0;      // all-gathered edge list (world > 1)
    spk_edge* d_wo_all = nullptr; uint64_t wo_all_cap = 0;    // gathered MI-without-gaps list (rank 0)
    spk_edge* d_edges_view = nullptr;                         // the list the last scan returned (d_edges, or d_merge after an exchange)
    unsigned long long* d_xcnt = nullptr;                     // [4 * world] doubles: per-rank {edges, wo, worst filter error, lost} (one all-gather)
    uint32_t* d_i8_colmax32_seed = nullptr;                   // seeded fp32 maxima kept for rescans (band / overflow retries)
    bool i8_seed_valid = false;
    Candidate* d_audit = nullptr; uint64_t audit_cap = 0;     // audit sample of the filter (pairs just below the candidate band)
    int i8_wr = 0;                         // weight rows the int8 operands / records were prepared for (2 = hi/lo, 1 = factorised)
    cudaStream_t copy_stream = nullptr;   // asynchronous result download (keep_on_device = 2)
    cudaEvent_t ev_ready = nullptr, ev_dl = nullptr;
    bool dl_pending = false;
    uint8_t* d_flags_buf = nullptr; uint64_t d_flags_cap = 0;     // spk_aracne's device flags (grow-only)
};

void resolve_timings(Ctx* c);               // call when the context's stream is idle

// ---- multi-GPU (spk_comm.cu): no-ops when c->world == 1; all calls are queued on c->stream
int comm_allreduce_sum_u64(Ctx* c, unsigned long long* d, size_t n);
int comm_allreduce_max_u32(Ctx* c, uint32_t* d, size_t n);
int comm_allreduce_max_u8(Ctx* c, uint8_t* d, size_t n);
int comm_allreduce_max_f64(Ctx* c, double* d, size_t n);
int comm_allreduce_sum_f64(Ctx* c, double* d, size_t n);
int comm_allgather_inplace(Ctx* c, void* d_buf, size_t bytes_per_rank);
int comm_allgatherv(Ctx* c, const void* d_mine, void* d_out, const size_t* bytes, const size_t* off);
int comm_gatherv_root(Ctx* c, const void* d_mine, void* d_out, const size_t* bytes, const size_t* off);
int comm_broadcast(Ctx* c, void* d_buf, size_t bytes, int root);
void comm_free(Ctx* c);

int fail(Ctx* c, int code, const char* fmt, ...);
int cuda_check(Ctx* c, cudaError_t e, const char* what);
#define SPK_CUDA(c, call) do { int _r = ::spk::cuda_check((c), (call), #call); if (_r) return _r; } while (0)

void* scratch(Ctx* c, size_t bytes);   // grows the context scratch buffer; nullptr on failure

// ---- prep (spk_prep.cu)
int prep_alignment(Ctx* c);                                   // colmajor, planes, masks from d_states
int prep_column_stats(Ctx* c, double pseudocount);            // d_marg, d_colstat, w_total
int prep_operands(Ctx* c, int n_terms);                       // d_U, d_V
int column_state_counts_device(Ctx* c, const uint8_t* d_raw, uint32_t n, uint64_t L, uint32_t* d_counts);   // [L][5] unweighted state counts
int gather_columns_device(Ctx* c, const uint8_t* d_raw, uint32_t n, uint64_t L, const uint32_t* d_idx, uint32_t Lk, uint8_t* d_out);

// ---- K1 (spk_hamming.cu)
int hamming_neighbour_counts(Ctx* c, uint64_t min_identical, int quirk, uint32_t part = 0, uint32_t n_parts = 1);

// ---- exact fp64 MI (spk_mi_exact.cu)
int mi_pairs_exact(Ctx* c, const uint32_t* d_pi, const uint32_t* d_pj, uint64_t npairs, double pc, double* d_mi);
// refine candidates -> d_refined (i,j,mi64), d_refined_wo (NaN when no gap column involved)
int mi_refine_prepare(Ctx* c, double pc);     // weight-grouped bit planes + entropy tables of the fp64 refine (asynchronous; idempotent until the weights change; world > 1: the plane build is sharded over the ranks and all-gathered)
// audit of the tensor-core filter: fp64 MI of the sample of pairs the seed pass emitted (mi32 > audit_lo): counters[8] = max |mi32 - mi64|
// (ordered bits), [9] = pairs with mi64 > thr that the main pass would NOT have refined (mi32 <= cand_lo), [10] = pairs audited
int mi_audit_append(Ctx* c, const Candidate* d_audit, const unsigned long long* d_n, uint64_t cap);      // audit sample -> marked entries of the candidate list
size_t mi_pairs_segmented_scratch(uint64_t total, uint32_t nseg);
int mi_pairs_segmented(Ctx* c, const uint32_t* d_pi, const uint32_t* d_pj, uint64_t total, uint64_t per, uint32_t nseg, double pc, double* d_mi, void* work);
int mi_refine_candidates(Ctx* c, uint64_t ncand, double pc, double wo_threshold, uint64_t* n_valid = nullptr);   // MI-without-gaps only when mi > wo_threshold
// exhaustive diagnostic engine: emits every pair of the partition as a candidate-free exact scan
int mi_scan_exact(Ctx* c, const spk_scan_params& p, uint64_t circle);
// threshold + colmax + compaction over refined candidates
int mi_compact(Ctx* c, uint64_t ncand, const spk_scan_params& p, uint64_t circle, float audit_cand_lo);
void mi_exact_free(Ctx* c);                                   // releases the weight-grouped bit planes

// ---- K2 (spk_mi_gemm.cu)
int mi_gemm_filter(Ctx* c, const spk_scan_params& p, int n_terms, double band, uint64_t circle);
int mi_gemm_debug_tile(Ctx* c, uint32_t ib, uint32_t jb, int n_terms, float* d_out);

// ---- K2, int8 compact engine (spk_mi_i8.cu)
int i8_build_layout(Ctx* c);                                   // after prep_alignment: classes, permutation
void i8_free(Ctx* c);
int mi_i8_filter(Ctx* c, const spk_scan_params& p, double band, uint64_t circle, int wr, bool reuse_seed);
int i8_default_wr();
int i8_tensor_peak(Ctx* c, uint32_t kblocks, uint32_t reps, float* ms, double* tops);   // MMA-only microbenchmark (roofline denominator)
int mi_i8_debug_tile(Ctx* c, uint32_t ti, uint32_t tj, double pc, uint32_t* d_out);

// ---- graph (spk_graph.cu)
int sort_edges_device(Ctx* c, spk_edge* d_edges, uint64_t n);
int merge_sorted_runs_device(Ctx* c, spk_edge* d_edges, const uint64_t* run_counts, uint32_t n_runs);
int aracne_device(Ctx* c, const spk_edge* d_edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, uint8_t* d_flags, bool by_node = false);

// ordered-int encodings for atomic max on floating point
__host__ __device__ inline uint32_t f32_ordered(float f) {
    uint32_t u;
#ifdef __CUDA_ARCH__
    u = __float_as_uint(f);
#else
    memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float f32_from_ordered(uint32_t u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
__host__ __device__ inline unsigned long long f64_ordered(double d) {
    unsigned long long u;
#ifdef __CUDA_ARCH__
    u = (unsigned long long)__double_as_longlong(d);
#else
    memcpy(&u, &d, 8);
#endif
    return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}
__host__ __device__ inline double f64_from_ordered(unsigned long long u) {
    u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double d; memcpy(&d, &u, 8); return d;
#endif
}

// distance used for the LD mask of colmax (apegrunt/Distance.hpp:28-37), size_t wrap included
__host__ __device__ inline unsigned long long genome_distance(unsigned long long p1, unsigned long long p2,
                                                              int linear, unsigned long long circle) {
    long long d0 = (long long)p1 - (long long)p2;
    unsigned long long d = (unsigned long long)(d0 < 0 ? -d0 : d0);
    if (linear) return d;
    unsigned long long alt = circle - d;
    return alt < d ? alt : d;
}

} // namespace spk

struct spk_ctx : public spk::Ctx {};
