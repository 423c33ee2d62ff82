// spk_mi_i8.cu -- K2, int8 "compact one-hot" engine: all-pairs weighted joint counts as a tcgen05
// kind::i8 GEMM (u8 x u8 -> s32, exact) over only the one-hot rows a column actually needs, with a
// fused fp32 MI / threshold / colmax-candidate / ballot-compaction epilogue.
//
// Replaces the same reference loops as spk_mi_gemm.cu
//   Weighted_crosstable_2Dblock::operator()   externals/apegrunt/include/misc/Crosstable.hpp:362-427
//   mutual_information_block_kernel::block     include/mi.hpp:108-181
//   MI_solver::operator() store loop           include/mi.hpp:414-461
// and is, like it, a FILTER in front of the fp64 refine (spk_mi_exact.cu).
//
// Two ideas on top of the dense bf16 formulation:
//  (1) Row compaction. A column with np present states needs only k = np-1 one-hot rows: the joint
//      table restricted to present states has (np_i x np_j) cells whose row / column sums are the
//      column marginals, so one row and one column (the highest-coded present state: the gap when the
//      column has gaps) are recovered by subtraction. SNP alignments are overwhelmingly biallelic
//      (k = 1 or 2), so the GEMM shrinks ~(4/1.8)^2 against the fixed 4-row layout. Columns are
//      permuted into classes of 1, 2 and 4 rows; an output tile is always 128 x 128 ROWS = (128/KI) x
//      (128/KJ) columns, KI <= KJ.
//  (2) Exact integer accumulation. Sample weights are quantised to 16 bits (w ~ q * delta,
//      q = 256*hi + lo); the N operand carries the two weight bytes as separate rows, the M operand is
//      the plain one-hot, and tcgen05.mma.kind::i8 accumulates in s32 -- at twice the bf16 rate and
//      without accumulation rounding. Marginals are integer sums of the same q, so the recovered
//      cells are exact for the quantised problem; the only filter error left is the weight
//      quantisation (<= delta/2 per sample) and the fp32 entropy arithmetic.
//
// GEMM (per tile):  A (M=128) rows = (i column, kept state a)          u8 one-hot
//                   B (N=256) rows = (j column, weight byte t, state b) u8 one-hot * {hi, lo}
//                   D[row][col] s32 in TMEM; N_ab = 256*D[.., t=0, b] + D[.., t=1, b].
//
// Kernel anatomy (persistent, 1 CTA/SM, 640 threads):
//   warp 0     TMA producer   cp.async.bulk.tensor 2D (128B swizzle), 4-stage ring of 48 KB (128 samples)
//   warp 1     MMA issuer     tcgen05.mma.cta_group::1.kind::i8 M128 N256 K32, commit -> mbarriers; owns TMEM
//   warps 2-17 epilogue       four warps per TMEM lane quadrant (each 64 of the 256 columns), free-running:
//                             tcgen05.ld 32x32b, hi/lo combine, KI-lane transpose, MI, candidates
//   warp 18    record stager  copies the next tile's column records into the other smem buffer (mbarrier handoff)
//   warp 19    tile scheduler global cursor -> (ti, tj), published through a 4-deep smem ring
// TMEM: 2 accumulator buffers x 256 columns, epilogue of tile t overlaps the mainloop of tile t+1.
#include <cuda.h>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include "spk_internal.h"
#include "spk_tcgen05.cuh"

namespace spk {

#ifndef SPK_I8_EPI_WARPS
#define SPK_I8_EPI_WARPS 16           // epilogue warps: 8 (two per TMEM lane quadrant) or 16 (four per quadrant)
#endif
constexpr int I8_EPI_WARPS = SPK_I8_EPI_WARPS;
constexpr int I8_EPI_THREADS = I8_EPI_WARPS * 32;
constexpr int I8_THREADS = 64 + I8_EPI_THREADS + 64;        // + TMA warp, MMA warp, record-staging warp, tile-scheduler warp
constexpr int I8_QDEPTH = 4;                             // dynamic tile scheduler ring depth
constexpr uint32_t I8_COLS_PER_WARP = 1024u / I8_EPI_WARPS;   // accumulator columns each epilogue warp drains (128 or 64)
constexpr uint32_t I8_BK = 128;                         // samples (= bytes) per stage row
constexpr uint32_t I8_A_BYTES = 128 * I8_BK;            // 16 KB
constexpr uint32_t I8_B_BYTES = 256 * I8_BK;            // 32 KB
constexpr uint32_t I8_STAGE_BYTES = I8_A_BYTES + I8_B_BYTES;
constexpr int I8_NSTAGES = 4;
#ifndef SPK_I8_SUPER
#define SPK_I8_SUPER 96
#endif
constexpr uint32_t I8_SUPER = SPK_I8_SUPER;             // i-tiles per super-row: their A panels (SUPER x 128 rows x n_pad B) stay L2-resident while B panels stream
constexpr uint32_t I8_IDESC =                           // cute::UMMA::InstrDescriptor
    (2u << 4) |                  // c_format = S32
    (0u << 7) | (0u << 10) |     // a_format = b_format = UINT8
    (0u << 15) | (0u << 16) |    // A, B K-major
    ((256u >> 3) << 17) |        // N = 256
    ((128u >> 4) << 24);         // M = 128

struct I8Params {
    const ColRec8* rec;
    Candidate* cand; unsigned long long cand_cap;
    unsigned long long* counters;          // [5] dynamic tile scheduler cursor, [6] experiment sink
    unsigned long long* cand_counter;      // slots reserved in `cand` (main pass: counters[0]; seed pass: the audit buffer's counter)
    unsigned long long* cand_overflow;     // set to 1 when a reservation runs past cand_cap (main pass: counters[4])
    uint32_t* colmax32;                    // [Lperm] running fp32 maxima (ordered-uint), permuted index
    const unsigned long long* sr_prefix;   // [n_sr+1] chunk prefix per super-row
    uint32_t n_sr, NT;
    uint32_t tile_start[3], col_start[3];
    uint32_t num_kb;                       // K blocks of 128 samples
    uint32_t part, n_parts;
    uint32_t pair;                         // 1: CTA-pair kernel (slot = two vertically adjacent tiles)
    uint32_t jt;                           // column tiles per B panel: 1 (hi/lo weight rows) or 2 (factorised weights, one row)
    uint32_t NTJ;                          // number of B panels = NT / jt
    const unsigned long long* seed_chunks; // seed pass: the sampled chunk ids (n_local_slots covers exactly these); nullptr = main pass
    unsigned long long n_local_slots;
    float pcq;                             // pseudocount in quantised-weight units
    float thr_lo, thr_hi, band;
    unsigned long long ld_threshold, circle; int linear;
    uint32_t* debug_out; uint32_t debug_ti, debug_tj;
};

// what a pair of a 1- or 2-row j class reads sits in the first 16 bytes (ONE LDS.128): mq[0..1], smq, cm_lo
struct __align__(16) JRec { uint32_t mq01[2]; uint32_t smq; float cm_lo; uint32_t mq23[2]; uint32_t k, orig;     // cm_lo: running max - band, +inf when closed
    __device__ __forceinline__ uint32_t mq(int b) const { return b < 2 ? mq01[b] : mq23[b - 2]; } };
template <int JMAX>                // JMAX = j columns a B panel can hold: 128 (two weight rows per j row) or 256 (one)
struct I8EpiSmemT {                // column records of one tile, staged by the stager warp one tile ahead
    JRec jrec[JMAX];
    unsigned long long jpos[JMAX];
    float jh[5 * (JMAX + 1)];       // H_j table by the i column's number of kept rows
    float ih2[2 * 128];             // H_i / 1/Z of the i columns for the two kept-row counts the j class can have
    float iinvz2[2 * 128];
    uint32_t cmj[JMAX];
    uint32_t imq[4 * 128];          // i columns, structure of arrays (lanes read consecutive columns)
    uint32_t imqd[128], ik[128], iorig[128], icm[128];
    unsigned long long ipos[128];
};
// kept-row counts possible in a j class: class 0 (1 row) {0,1}, class 1 (2 rows) {2}, class 2 (4 rows) {3,4}
__device__ __forceinline__ uint32_t i8_class_klo(uint32_t c) { return c == 0u ? 0u : (c == 1u ? 2u : 3u); }
__device__ __forceinline__ uint32_t i8_class_khi(uint32_t c) { return c == 0u ? 1u : (c == 1u ? 2u : 4u); }

// local slot -> (ti, tj); SUPER consecutive slots share the B (j) panel
__device__ __forceinline__ bool i8_slot_to_tile(const I8Params& p, unsigned long long ls, uint32_t& ti, uint32_t& tj) {
    if (p.debug_out) { ti = p.debug_ti; tj = p.debug_tj / p.jt; return ls == 0; }
    const uint32_t spc = p.pair ? I8_SUPER / 2 : I8_SUPER;            // slots per chunk (a CTA pair takes two i-tiles per slot)
    const unsigned long long q = p.seed_chunks ? p.seed_chunks[ls / spc] : (ls / spc) * p.n_parts + p.part;
    const uint32_t w = (uint32_t)(ls % spc) << p.pair;
    if (q >= p.sr_prefix[p.n_sr]) return false;
    uint32_t lo = 0, hi = p.n_sr;
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (p.sr_prefix[mid] <= q) lo = mid; else hi = mid; }
    ti = lo * I8_SUPER + w;
    tj = lo * (I8_SUPER / p.jt) + (uint32_t)(q - p.sr_prefix[lo]);           // B panel index: column tiles jt*tj .. jt*tj + jt-1
    return ti < p.NT && tj < p.NTJ && ti <= p.jt * tj + (p.jt - 1u);
}
__device__ __forceinline__ uint32_t i8_tile_class(const I8Params& p, uint32_t t) {
    return (t >= p.tile_start[2]) ? 2u : (t >= p.tile_start[1]) ? 1u : 0u;
}

template <int W> __device__ __forceinline__ void i8_tmem_ld(uint32_t taddr, uint32_t* r);
template <> __device__ __forceinline__ void i8_tmem_ld<8>(uint32_t taddr, uint32_t* r) { tmem_ld8_u32(taddr, r); tmem_ld_wait(); }
template <> __device__ __forceinline__ void i8_tmem_ld<16>(uint32_t taddr, uint32_t* r) { tmem_ld16_u32(taddr, r); tmem_ld_wait(); }
template <> __device__ __forceinline__ void i8_tmem_ld<32>(uint32_t taddr, uint32_t* r) {
    tmem_ld16_u32(taddr, r); tmem_ld16_u32(taddr + 16u, r + 16); tmem_ld_wait();
}

// colmax bookkeeping, entered only by lanes whose pair is within `band` of the running maximum of one of
// its two columns while that maximum has not yet cleared thr+band (rare after the first tiles of a column:
// closed columns carry +inf in cm_lo). Lanes that arrive together and share the j column reduce their
// maximum with match/redux so a burst emits ~one candidate per warp, not one per lane.
struct I8OpenRes { float cm_i, cm_i_lo; uint32_t emit; };
template <int JMAX>
__device__ __noinline__ I8OpenRes i8_open_columns_path(const I8Params& p, I8EpiSmemT<JMAX>& sm, float mi32, uint32_t i_local,
                                                       uint32_t j_local, uint32_t gcol_j, float cm_i, float cm_i_lo) {
    I8OpenRes res; res.cm_i = cm_i; res.cm_i_lo = cm_i_lo; res.emit = 0u;
    const unsigned long long pos_j = sm.jpos[j_local];
    if (!(genome_distance(sm.ipos[i_local], pos_j, p.linear, p.circle) > p.ld_threshold)) return res;
    bool emit = false;
    if (mi32 >= cm_i_lo) {                                         // i side: lane-local running maximum
        const float c = fmaxf(cm_i, mi32);
        res.cm_i = c;
        if (c > p.thr_hi) res.cm_i_lo = INFINITY; else { res.cm_i_lo = c - p.band; emit = true; }
    }
    const uint32_t peers = __match_any_sync(__activemask(), j_local);
    const uint32_t mine = f32_ordered(mi32);
    const uint32_t gmax = __reduce_max_sync(peers, mine);
    const uint32_t cur = sm.cmj[j_local];
    const float nm = f32_from_ordered(gmax > cur ? gmax : cur);
    if (gmax > cur && (peers & ((1u << (threadIdx.x & 31u)) - 1u)) == 0u) {       // leader of the peer group
        atomicMax(&sm.cmj[j_local], gmax);
        atomicMax(&p.colmax32[gcol_j], gmax);
        sm.jrec[j_local].cm_lo = (nm > p.thr_hi) ? INFINITY : nm - p.band;        // benign race: any stale value is a lower bound
    }
    if (!(nm > p.thr_hi) && mi32 >= nm - p.band) emit = true;
    res.emit = emit ? 1u : 0u;
    return res;
}

// ---------------------------------------------------------------------------------------------
// candidate emission: every epilogue warp owns a chunk of I8_CH slots of the candidate array (one atomicAdd per
// chunk, the next chunk is reserved ahead of need), so a warp that finds candidates does not sit out a global
// atomic round trip (~1 us) per ballot -- measured: tests + emission cost 6 ms of a 55 ms launch that way.
// Unused slots of a chunk are filled with an invalid marker (i = j = I8_INVALID) which the refine pass drops.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t I8_CH = 128;
struct I8Emit { uint32_t cur, used, next; };      // cur: chunk id (warp-uniform; I8_INVALID = none yet), next: reserved ahead (lane 0 only)

__device__ __forceinline__ uint32_t i8_reserve_chunk(const I8Params& p) {
    return (uint32_t)(atomicAdd(p.cand_counter, (unsigned long long)I8_CH) / I8_CH);
}
__device__ __forceinline__ void i8_fill_invalid(const I8Params& p, uint32_t chunk, uint32_t from, uint32_t lane) {
    Candidate cd; cd.i = I8_INVALID; cd.j = I8_INVALID; cd.mi32 = 0.f; cd.pad = 0;
    for (uint32_t k = from + lane; k < I8_CH; k += 32) {
        const unsigned long long idx = (unsigned long long)chunk * I8_CH + k;
        if (idx < p.cand_cap) p.cand[idx] = cd;
    }
}
__device__ __noinline__ void i8_emit_switch(const I8Params& p, I8Emit& em, uint32_t lane) {
    if (em.cur != I8_INVALID) i8_fill_invalid(p, em.cur, em.used, lane);
    uint32_t nx = em.next;
    if (lane == 0 && nx == I8_INVALID) nx = i8_reserve_chunk(p);
    nx = __shfl_sync(0xffffffffu, nx, 0);
    em.cur = nx; em.used = 0; em.next = I8_INVALID;
    if ((unsigned long long)(nx + 1u) * I8_CH > p.cand_cap && lane == 0) *p.cand_overflow = 1ull;
}
__device__ __forceinline__ void i8_emit_flush(const I8Params& p, I8Emit& em, uint32_t lane) {
    if (em.cur != I8_INVALID) i8_fill_invalid(p, em.cur, em.used, lane);
    const uint32_t nx = __shfl_sync(0xffffffffu, em.next, 0);
    if (nx != I8_INVALID) i8_fill_invalid(p, nx, 0, lane);
}

// ---------------------------------------------------------------------------------------------
// epilogue of one tile for the class pair (KI rows per i column, KJ rows per j column), KI <= KJ
// ---------------------------------------------------------------------------------------------
// WR = weight rows per j row in the N operand: 2 = {hi, lo} bytes of a 16-bit weight, 1 = one byte b_s (A carries a_s)
template <int KI, int KJ, int WR>
__device__ __forceinline__ void i8_epilogue_tile(const I8Params& p, I8EpiSmemT<256 / WR>& sm, I8Emit& em, uint32_t tmem_acc, uint32_t q, uint32_t half,
                                                 uint32_t lane, bool diag, uint32_t ci0, uint32_t cj0) {
    constexpr int LOADW = ((KI == 4) ? 32 : 16) / (3 - WR);              // TMEM columns per load; WR = 1: half the columns hold the same number of j rows
    constexpr int NJ = LOADW / (WR * KJ);              // j columns per load
    constexpr int PPL = NJ / KI;                       // pairs per lane per load
    constexpr int JMAX = 256 / WR;
    constexpr int WJT = 128 / KJ;                      // j columns of one 128-row column tile (debug output geometry)
    static_assert(PPL >= 1, "tile geometry");
    const uint32_t row = q * 32u + lane;
    const uint32_t i_local = row / KI, g = lane % KI;
    const uint32_t pi = ci0 + i_local;
    // per-lane constants of the i column (staged in shared memory by the stager warp)
    uint32_t mi_[KI];
#pragma unroll
    for (int a = 0; a < KI; ++a) mi_[a] = sm.imq[a * 128 + i_local];
    // (position and original index stay in shared memory: they are only needed on the rare open-column / emission paths)
    const uint32_t mi_d = sm.imqd[i_local], ki = sm.ik[i_local];
    const bool vi = sm.iorig[i_local] != I8_INVALID;
    const float cm_i0 = f32_from_ordered(sm.icm[i_local]);
    float cm_i = cm_i0;
    float cm_i_lo = (vi && !(cm_i0 > p.thr_hi)) ? cm_i0 - p.band : INFINITY;      // +inf: column closed (covered by the threshold rule)
    constexpr uint32_t KJ_HI = (KJ == 1) ? 1u : (KJ == 2 ? 2u : 4u);              // j columns of this class keep KJ_HI or (KJ != 2) one row fewer
    const float invZ_lo = sm.iinvz2[i_local], invZ_hi = sm.iinvz2[128 + i_local];
    const float ih_lo = sm.ih2[i_local], ih_hi = sm.ih2[128 + i_local];
    const float* jh_row = sm.jh + ki * (uint32_t)(JMAX + 1);
    const float ln2 = 0.69314718055994530942f;
    const float thr_lo = p.thr_lo;

#pragma unroll 1
    for (uint32_t ld = 0; ld < I8_COLS_PER_WARP / LOADW; ++ld) {
        const uint32_t col0 = half * I8_COLS_PER_WARP + ld * LOADW;
        uint32_t r[LOADW];
        i8_tmem_ld<LOADW>(tmem_acc + ((q * 32u) << 16) + col0, r);
#if defined(SPK_I8_EPI_LEVEL) && SPK_I8_EPI_LEVEL == 1          // experiment: TMEM loads only
        { uint32_t x = 0;
#pragma unroll
          for (int k = 0; k < LOADW; ++k) x ^= r[k];
          if (x == 0xdeadbeefu) p.counters[6] = x;
          continue; }
#endif
        uint32_t v[NJ * KJ];
#pragma unroll
        for (int jj = 0; jj < NJ; ++jj)
#pragma unroll
            for (int b = 0; b < KJ; ++b)
                v[jj * KJ + b] = (WR == 2) ? (r[jj * 2 * KJ + b] << 8) + r[jj * 2 * KJ + KJ + b] : r[jj * KJ + b];
        // transpose inside the KI-lane group: afterwards v[(s*KI + a)*KJ + b] = N[a][b] of j = s*KI + g
        if (KI == 2) {
#pragma unroll
            for (int s = 0; s < NJ / 2; ++s)
#pragma unroll
                for (int b = 0; b < KJ; ++b) {
                    const int s0 = (2 * s) * KJ + b, s1 = (2 * s + 1) * KJ + b;
                    const uint32_t send = g ? v[s0] : v[s1];
                    const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 1);
                    if (g) v[s0] = recv; else v[s1] = recv;
                }
        } else if (KI == 4) {
            const bool b0 = lane & 1u, b1 = lane & 2u;
#pragma unroll
            for (int s = 0; s < NJ / 4; ++s) {
#pragma unroll
                for (int pr = 0; pr < 2; ++pr)
#pragma unroll
                    for (int b = 0; b < KJ; ++b) {
                        const int s0 = (4 * s + 2 * pr) * KJ + b, s1 = (4 * s + 2 * pr + 1) * KJ + b;
                        const uint32_t send = b0 ? v[s0] : v[s1];
                        const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 1);
                        if (b0) v[s0] = recv; else v[s1] = recv;
                    }
#pragma unroll
                for (int pr = 0; pr < 2; ++pr)
#pragma unroll
                    for (int b = 0; b < KJ; ++b) {
                        const int s0 = (4 * s + pr) * KJ + b, s1 = (4 * s + pr + 2) * KJ + b;
                        const uint32_t send = b1 ? v[s0] : v[s1];
                        const uint32_t recv = __shfl_xor_sync(0xffffffffu, send, 2);
                        if (b1) v[s0] = recv; else v[s1] = recv;
                    }
            }
        }
        if (p.debug_out) {                                     // tests only: raw counts of the requested tile (one branch per load)
#pragma unroll
            for (int s = 0; s < PPL; ++s) {
                const uint32_t j_local = col0 / (WR * KJ) + s * KI + g;
                if (j_local / WJT == p.debug_tj % p.jt) {
#pragma unroll
                    for (int a = 0; a < KI; ++a)
#pragma unroll
                        for (int b = 0; b < KJ; ++b)
                            p.debug_out[(((size_t)i_local * WJT + j_local % WJT) * KI + a) * KJ + b] = v[(s * KI + a) * KJ + b];
                }
            }
            continue;
        }
        // phase 1: branch-free MI of the PPL pairs of this load (independent chains -> MUFU / LDS latencies overlap). No validity
        // logic here: padding / below-diagonal pairs are dropped in phase 2, which only the rare load with a passing pair enters
        // (the per-pair valid-select, NaN constant and debug branch had cost 10 of 52 instructions per pair).
        float mi_arr[PPL];
        bool need = false;
        const float t_i = fminf(thr_lo, cm_i_lo);                  // cm_i_lo only moves in phase 2
#pragma unroll
        for (int s = 0; s < PPL; ++s) {
            const uint32_t j_local = col0 / (WR * KJ) + s * KI + g;
            const JRec jr = sm.jrec[j_local];
            const uint32_t kj = jr.k;
            const bool khi = (KJ == 2) || (kj == KJ_HI);
            const float invZ = khi ? invZ_hi : invZ_lo;
            const float hsum = (khi ? ih_hi : ih_lo) + jh_row[j_local];
            const float pcZ = p.pcq * invZ;
            // p >= pc/Z is far above the denormal range: plain MUFU.LG2 (ftz) without the denormal rescue sequence
            auto xlg = [&](uint32_t nq) {
                const float pr = fmaf((float)nq, invZ, pcZ);
                float l; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(pr));
                return pr * l;
            };
            // integer pass: row sums, column sums, total of the kept cells
            uint32_t colsum[KJ], rs[KI];
#pragma unroll
            for (int b = 0; b < KJ; ++b) colsum[b] = 0;
            uint32_t tot = 0;
#pragma unroll
            for (int a = 0; a < KI; ++a) {
                rs[a] = 0;
#pragma unroll
                for (int b = 0; b < KJ; ++b) { const uint32_t nab = v[(s * KI + a) * KJ + b]; rs[a] += nab; colsum[b] += nab; }
                tot += rs[a];
            }
            // S = sum of p log2 p over the cells of the (ki+1) x (kj+1) table: the dropped-i row starts the chain, every kept row
            // continues it (one FFMA per cell; a row / column the column does not have is skipped with a select, never added as 0 + x)
            float S = xlg(mi_d - jr.smq + tot);                          // cell (dropped_i, dropped_j)
#pragma unroll
            for (int b = 0; b < KJ; ++b) {
                const bool cv = (KJ == 2) ? true : (KJ == 1 ? (kj >= 1u) : (b < 3 || kj == 4u));
                const float t = S + xlg(jr.mq(b) - colsum[b]);           // cell (dropped_i, b)
                S = cv ? t : S;
            }
#pragma unroll
            for (int a = 0; a < KI; ++a) {
                const bool rv = (KI == 2) ? true : (KI == 1 ? (ki >= 1u) : (a < 3 || ki == 4u));
                float racc = S + xlg(mi_[a] - rs[a]);                    // cell (a, dropped_j)
#pragma unroll
                for (int b = 0; b < KJ; ++b) {
                    const bool cv = (KJ == 2) ? true : (KJ == 1 ? (kj >= 1u) : (b < 3 || kj == 4u));
                    const float t = racc + xlg(v[(s * KI + a) * KJ + b]);
                    racc = cv ? t : racc;
                }
                S = rv ? racc : S;
            }
            mi_arr[s] = fmaf(S, ln2, -hsum);
            need = need || (mi_arr[s] >= fminf(t_i, jr.cm_lo));          // superset of the exact tests below
        }
        if (!__any_sync(0xffffffffu, need)) continue;             // nothing to test or emit in this load (the common case)
        // phase 2: validity, threshold / open-column tests and candidate emission
#pragma unroll
        for (int s = 0; s < PPL; ++s) {
            const uint32_t j_local = col0 / (WR * KJ) + s * KI + g;
            const uint32_t pj = cj0 + j_local;
            const bool valid = vi && (sm.jrec[j_local].orig != I8_INVALID) && (!diag || pi < pj);
            const float mi32 = valid ? mi_arr[s] : __int_as_float(0x7fc00000);     // NaN fails every test below
            bool emit;
            if (p.seed_chunks) {
                // seed pass: the running LD-admissible maxima of both columns; the only pairs it emits are the AUDIT sample of
                // the filter (mi32 > thr - 4*band, into the audit buffer: see mi_audit_*_kernel, spk_mi_exact.cu)
                const bool adm = (mi32 == mi32) && genome_distance(sm.ipos[i_local], sm.jpos[j_local], p.linear, p.circle) > p.ld_threshold;
                const float m = adm ? mi32 : -INFINITY;
                cm_i = fmaxf(cm_i, m);
                constexpr uint32_t GM = (KI == 1) ? 0xffffffffu : (KI == 2 ? 0x55555555u : 0x11111111u);   // lanes that share this j column
                const uint32_t o = f32_ordered(m);
                const uint32_t gmax = __reduce_max_sync(GM << g, o);
                if (lane == g && gmax > sm.cmj[j_local]) {
                    if (atomicMax(&sm.cmj[j_local], gmax) < gmax) atomicMax(&p.colmax32[pj], gmax);
                }
                emit = mi32 > p.thr_lo;                                        // thr_lo of the seed launch = thr - 4*band (+inf: no audit)
            } else {
            const float cm_lo_j = sm.jrec[j_local].cm_lo;
#if defined(SPK_I8_EPI_LEVEL) && SPK_I8_EPI_LEVEL == 2          // experiment: MI arithmetic without tests / emission
            if (mi32 == 123.456f) p.counters[6] = 1ull;
            continue;
#endif
            emit = mi32 > p.thr_lo;                                        // false for NaN
#if defined(SPK_I8_EPI_LEVEL) && SPK_I8_EPI_LEVEL == 3          // experiment: no open-column path
            if (mi32 == 123.456f) {
#else
            if (mi32 >= fminf(cm_i_lo, cm_lo_j)) {                         // false for NaN (invalid pair)
#endif
                const I8OpenRes orr = i8_open_columns_path(p, sm, mi32, i_local, j_local, pj, cm_i, cm_i_lo);
                cm_i = orr.cm_i; cm_i_lo = orr.cm_i_lo; emit = emit || (orr.emit != 0u);
            }
#if defined(SPK_I8_EPI_LEVEL) && SPK_I8_EPI_LEVEL == 4          // experiment: no emission
            if (emit && mi32 == 123.456f) p.counters[6] = 1ull;
            continue;
#endif
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, emit);
            if (bal) {
                const uint32_t nb = __popc(bal);
                if (em.cur == I8_INVALID || em.used + nb > I8_CH) i8_emit_switch(p, em, lane);
                if (emit) {
                    const unsigned long long idx = (unsigned long long)em.cur * I8_CH + em.used + __popc(bal & ((1u << lane) - 1u));
                    if (idx < p.cand_cap) {
                        Candidate cd;
                        const uint32_t orig_i = sm.iorig[i_local], orig_j = sm.jrec[j_local].orig;
                        cd.i = min(orig_i, orig_j); cd.j = max(orig_i, orig_j); cd.mi32 = mi32; cd.pad = 0;
                        p.cand[idx] = cd;
                    }
                }
                em.used += nb;
                if (lane == 0 && em.next == I8_INVALID && em.used > I8_CH / 2) em.next = i8_reserve_chunk(p);   // result is not needed before the switch
            }
        }
    }
    if (!p.debug_out && vi && cm_i > cm_i0) atomicMax(&p.colmax32[pi], f32_ordered(cm_i));
}

// ---------------------------------------------------------------------------------------------
// NCTA = 1: one CTA per tile (tcgen05 cta_group::1, M128 N256). NCTA = 2: a CTA pair (cluster of 2 on one TPC)
// works on two vertically adjacent tiles (2x, tj), (2x+1, tj) with ONE cta_group::2 MMA stream (M256 N256)
// issued by the pair's leader: each CTA stages its own A panel and only HALF of the shared B panel, so the
// L2 -> SM operand traffic per tile falls from 48 KB to 32 KB per 128 samples. (ncu on the 1-CTA kernel:
// lts2xbar 71 % busy at 69 % tensor-pipe activity -- the L2 feed, not the pipe, was the limiter.)
// WR = 2: the N operand carries the {hi, lo} bytes of a 16-bit weight as two rows per j row (a B panel = one column tile);
// WR = 1: factorised weights q_s = a_s * b_s, A carries a_s, B carries b_s in ONE row per j row, so a 256-row B panel
// holds TWO column tiles and every MMA covers twice the pairs (the record buffers double, one operand stage goes).
template <int NCTA, int WR> struct I8Cfg {
    static constexpr uint32_t B_BYTES = (NCTA == 2 ? 128u : 256u) * I8_BK, STAGE = I8_A_BYTES + B_BYTES;
    static constexpr int NSTAGES = (NCTA == 2) ? (WR == 2 ? 6 : 5) : (WR == 2 ? 4 : 3);
};

template <int NCTA, int WR>
__global__ void __launch_bounds__(I8_THREADS, 1)
mi_i8_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ I8Params p) {
    using Cfg = I8Cfg<NCTA, WR>;
    using I8EpiSmem = I8EpiSmemT<256 / WR>;
    constexpr uint32_t JT = 2u / WR;                          // column tiles per B panel
    constexpr bool PAIR = (NCTA == 2);
    constexpr int NST = Cfg::NSTAGES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // The dynamic shared segment is the only shared object of this kernel and is declared 1024-aligned, so every
    // buffer below sits at a compile-time offset from the symbol (LDS/STS with immediate offsets; a run-time
    // alignment fix-up made the compiler rebuild the shared-window base in front of every access).
    uint8_t* smem = smem_raw;
    if (threadIdx.x == 0 && (smem_u32(smem_raw) & 1023u) != 0u) { printf("spk: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    uint8_t* tiles = smem;
    constexpr size_t SM_STRIDE = (sizeof(I8EpiSmem) + 15) & ~(size_t)15;
    uint8_t* sm_base = smem + (size_t)NST * Cfg::STAGE;                        // two record buffers (tile t, tile t+1)
    uint64_t* full_bar = (uint64_t*)(sm_base + 2 * SM_STRIDE);                 // pair: only the leader's are used
    uint64_t* empty_bar = full_bar + NST;
    uint64_t* tfull_bar = empty_bar + NST;               // [2]
    uint64_t* tempty_bar = tfull_bar + 2;                // [2] pair: only the leader's, both CTAs' epilogues arrive
    uint64_t* sfull_bar = tempty_bar + 2;                // [2] column records of a tile staged
    uint64_t* sempty_bar = sfull_bar + 2;                // [2] ... and released by all epilogue warps
    uint64_t* qfull_bar = sempty_bar + 2;                // [I8_QDEPTH] tile scheduler ring: entry published
    uint64_t* qempty_bar = qfull_bar + I8_QDEPTH;        // [I8_QDEPTH] ... and read by every consumer role (pair: leader's)
    uint2* q_tile = (uint2*)(qempty_bar + I8_QDEPTH);    // [I8_QDEPTH] (ti, tj); ti = I8_INVALID ends the stream
    uint32_t* tmem_slot = (uint32_t*)(q_tile + I8_QDEPTH);

    const uint32_t warp = __shfl_sync(0xffffffffu, threadIdx.x / 32, 0), lane = threadIdx.x % 32;     // warp-uniform for the compiler
    const uint32_t rank = PAIR ? cluster_ctarank() : 0u;
    const bool leader = (rank == 0u);

    if (threadIdx.x == 0) {
        for (int s = 0; s < NST; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], NCTA * I8_EPI_WARPS);
            mbar_init(&sfull_bar[b], 1); mbar_init(&sempty_bar[b], I8_EPI_WARPS);
        }
        for (int r = 0; r < I8_QDEPTH; ++r) { mbar_init(&qfull_bar[r], 1); mbar_init(&qempty_bar[r], NCTA * (I8_EPI_WARPS + 2) + 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1) {
        if constexpr (PAIR) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tcgen05_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // queue helpers: entries are written by the leader's scheduler (into both CTAs), released to the leader's barrier
    auto q_wait = [&](uint32_t r, uint32_t rph) { if constexpr (PAIR) mbar_wait_cluster(&qfull_bar[r], rph); else mbar_wait(&qfull_bar[r], rph); };
    auto q_release = [&](uint32_t r) { if (PAIR && !leader) mbar_arrive_remote_nofence(&qempty_bar[r], 0); else mbar_arrive(&qempty_bar[r]); };
    // does THIS CTA have a tile to finish for queue entry (t0, tj)? (pair: the odd CTA idles below the diagonal / past the last tile)
    auto own_valid = [&](uint32_t ti_own, uint32_t tj) {
        if constexpr (!PAIR) return true;
        return ti_own < p.NT && ti_own <= JT * tj + (JT - 1u) && (!p.debug_out || ti_own == p.debug_ti);
    };

    if (warp == 3 + I8_EPI_WARPS) {
        // ===================== dynamic tile scheduler (leader CTA of a pair; every CTA otherwise) =====================
        // Tiles are handed out from ONE global cursor in slot order, so the 148 CTAs always work inside a window of
        // ~148 consecutive slots (a few B panels + one super-row of A panels): the operand working set stays in L2
        // however unevenly the per-class epilogues run. (A static round-robin lets CTAs drift apart until B panels are
        // evicted between their users: measured 630 GB of DRAM reads for 1.7 GB of operands.) The cursor round trip and
        // the slot -> tile search (~2 us of dependent global accesses) run up to I8_QDEPTH tiles ahead in their own warp;
        // done in the TMA thread they drained the 4-stage operand ring between tiles (tensor pipe 69 % active).
        if (lane == 0 && leader) {
            uint32_t r = 0, rph = 0;
            bool debug_done = false;
            while (true) {
                uint32_t ti = I8_INVALID, tj = 0;
                if (p.debug_out) { if (!debug_done) { ti = PAIR ? (p.debug_ti & ~1u) : p.debug_ti; tj = p.debug_tj / JT; debug_done = true; } }
                else {
                    while (true) {
                        const unsigned long long ls = atomicAdd(&p.counters[5], 1ull);
                        if (ls >= p.n_local_slots) { ti = I8_INVALID; break; }
                        if (i8_slot_to_tile(p, ls, ti, tj)) break;
                        ti = I8_INVALID;
                    }
                }
                if constexpr (PAIR) mbar_wait_cluster(&qempty_bar[r], rph ^ 1u); else mbar_wait(&qempty_bar[r], rph ^ 1u);
                q_tile[r] = make_uint2(ti, tj);
                if constexpr (PAIR) {
                    st_remote_u64(&q_tile[r], 1, ((unsigned long long)tj << 32) | ti);
                    mbar_arrive_remote(&qfull_bar[r], 1);
                }
                mbar_arrive(&qfull_bar[r]);
                if (++r == I8_QDEPTH) { r = 0; rph ^= 1u; }
                if (ti == I8_INVALID) break;
            }
        }
    } else if (warp == 0) {
        // ===================== TMA producer (whole warp converged, one elected lane issues) =====================
        uint32_t stage = 0, phase = 0, r = 0, rph = 0;
        while (true) {
            q_wait(r, rph);
            const uint2 qt = q_tile[r];
            const uint32_t ti = __shfl_sync(0xffffffffu, qt.x, 0), tj = qt.y;      // the shuffle consumes the load before the slot is handed back
            if (elect_one()) q_release(r);
            if (++r == I8_QDEPTH) { r = 0; rph ^= 1u; }
            if (ti == I8_INVALID) break;
            const int32_t ya = (int32_t)((ti + rank) * 128u);
            const int32_t yb = (int32_t)(PAIR ? tj * 256u + rank * 128u : tj * 256u);
            for (uint32_t kb = 0; kb < p.num_kb; ++kb) {
                mbar_wait(&empty_bar[stage], phase ^ 1u);
                uint8_t* st = tiles + (size_t)stage * Cfg::STAGE;
                if (elect_one()) {
                    if constexpr (PAIR) {
                        if (leader) mbar_expect_tx(&full_bar[stage], 2u * Cfg::STAGE);       // both CTAs' boxes land on the leader's barrier
                        tma_load_2d_pair(&map_b, &full_bar[stage], st, (int32_t)(kb * I8_BK), yb);
                        tma_load_2d_pair(&map_a, &full_bar[stage], st + Cfg::B_BYTES, (int32_t)(kb * I8_BK), ya);
                    } else {
                        mbar_expect_tx(&full_bar[stage], Cfg::STAGE);
                        tma_load_2d(&map_b, &full_bar[stage], st, (int32_t)(kb * I8_BK), yb);
                        tma_load_2d(&map_a, &full_bar[stage], st + Cfg::B_BYTES, (int32_t)(kb * I8_BK), ya);
                    }
                }
                __syncwarp();
                if (++stage == NST) { stage = 0; phase ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (pair: the leader's warp drives both tensor cores) =====================
        // The warp stays converged and one elected lane issues: with an `if (lane == 0)` region the compiler wraps
        // every tcgen05 instruction in a uniform-register waterfall loop (~90 instructions per 128 samples, as
        // long as the four MMAs themselves take -- the issue thread, not the pipe, was the bottleneck).
        if (leader) {
            constexpr uint32_t IDESC = PAIR ? ((I8_IDESC & ~(31u << 24)) | ((256u >> 4) << 24)) : I8_IDESC;
            uint32_t stage = 0, phase = 0, it = 0, r = 0, rph = 0;
            // shared-memory matrix descriptors: the high word is a constant and the 14-bit address field of the low word never
            // carries (shared memory < 256 KB), so a stage's descriptors are one 32-bit add away from the previous stage's and the
            // four K steps of a stage are compile-time offsets (the 64-bit descriptor arithmetic per K block was ~25 of the ~70
            // dependent uniform instructions the issuing thread spends per 128 samples)
            constexpr uint32_t DESC_HI = 64u | (1u << 14) | (2u << 29);          // SBO = 1024 B >> 4, version 1, SWIZZLE_128B
            const uint32_t desc_lo0 = ((smem_u32(tiles) >> 4) & 0x3FFFu) | (1u << 16);
            uint32_t desc_lo = desc_lo0;                                         // B operand of the current stage; A = + B_BYTES >> 4
            while (true) {
                q_wait(r, rph);
                const uint32_t ti = __shfl_sync(0xffffffffu, q_tile[r].x, 0);
                if (elect_one()) q_release(r);
                if (++r == I8_QDEPTH) { r = 0; rph ^= 1u; }
                if (ti == I8_INVALID) break;
                const uint32_t buf = it & 1u, acc_phase = (it >> 1) & 1u;
                if constexpr (PAIR) mbar_wait_cluster(&tempty_bar[buf], acc_phase ^ 1u); else mbar_wait(&tempty_bar[buf], acc_phase ^ 1u);
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + buf * 256u;
                for (uint32_t kb = 0; kb < p.num_kb; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t lb = desc_lo, la = desc_lo + (Cfg::B_BYTES >> 4);
                    if (elect_one()) {
#pragma unroll
                        for (uint32_t k = 0; k < I8_BK / 32; ++k) {      // +32 B along K = +2 in the descriptor's 16-byte address field
                            const uint64_t da = ((uint64_t)DESC_HI << 32) | (la + 2u * k), db = ((uint64_t)DESC_HI << 32) | (lb + 2u * k);
                            if constexpr (PAIR) umma_i8_pair(d_tmem, da, db, IDESC, (kb | k) ? 1u : 0u);
                            else umma_i8(d_tmem, da, db, IDESC, (kb | k) ? 1u : 0u);
                        }
                        if constexpr (PAIR) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
                    }
                    __syncwarp();
                    desc_lo += Cfg::STAGE >> 4;
                    if (++stage == NST) { stage = 0; phase ^= 1u; desc_lo = desc_lo0; }
                }
                if (elect_one()) { if constexpr (PAIR) umma_commit_pair(&tfull_bar[buf]); else umma_commit(&tfull_bar[buf]); }
                __syncwarp();
                ++it;
            }
        }
    } else if (warp == 2 + I8_EPI_WARPS) {
        // ===================== record stager: runs one tile ahead of the epilogue =====================
        // Copies the column records of the next tile from global memory into the other smem buffer, so the
        // epilogue warps never expose that latency and need no CTA-wide barrier between tiles.
        uint32_t it = 0, r = 0, rph = 0;
        while (true) {
            q_wait(r, rph);
            const uint2 qt = q_tile[r];
            const uint32_t t0 = __shfl_sync(0xffffffffu, qt.x, 0);                 // consumes the load before the slot is handed back
            if (lane == 0) q_release(r);
            if (++r == I8_QDEPTH) { r = 0; rph ^= 1u; }
            if (t0 == I8_INVALID) break;
            const uint32_t ti = t0 + rank, tj = qt.y;
            if (!own_valid(ti, tj)) continue;
            const uint32_t buf = it & 1u, ph = (it >> 1) & 1u;
            I8EpiSmem& sm = *(I8EpiSmem*)(sm_base + buf * SM_STRIDE);
            const uint32_t ci = i8_tile_class(p, ti), cj = i8_tile_class(p, JT * tj);   // (class tile counts are even: a panel never straddles classes)
            const uint32_t WI = 128u >> ci, WJ = JT * (128u >> cj);
            const uint32_t ci0 = p.col_start[ci] + (ti - p.tile_start[ci]) * WI;
            const uint32_t cj0 = p.col_start[cj] + (JT * tj - p.tile_start[cj]) * (128u >> cj);
            mbar_wait(&sempty_bar[buf], ph ^ 1u);
            for (uint32_t e = lane; e < WJ; e += 32) {
                const ColRec8 rc = p.rec[cj0 + e];
                JRec jr;
                jr.mq01[0] = rc.mq[0]; jr.mq01[1] = rc.mq[1]; jr.mq23[0] = rc.mq[2]; jr.mq23[1] = rc.mq[3];
                jr.smq = rc.smq; jr.k = rc.k; jr.orig = rc.orig;
                const uint32_t cm = p.colmax32[cj0 + e];
                const float cmf = f32_from_ordered(cm);
                jr.cm_lo = (rc.orig != I8_INVALID && !(cmf > p.thr_hi)) ? cmf - p.band : INFINITY;
                sm.jrec[e] = jr;
                sm.jpos[e] = ((unsigned long long)rc.pos_hi << 32) | rc.pos_lo;
#pragma unroll
                for (int m = 0; m < 5; ++m) sm.jh[m * (256 / WR + 1) + e] = rc.h[m];
                sm.cmj[e] = cm;
            }
            const uint32_t klo = i8_class_klo(cj), khi = i8_class_khi(cj);
            for (uint32_t e = lane; e < WI; e += 32) {
                const ColRec8* rc = p.rec + ci0 + e;
                sm.ih2[e] = rc->h[klo]; sm.ih2[128 + e] = rc->h[khi];
                sm.iinvz2[e] = rc->invz[klo]; sm.iinvz2[128 + e] = rc->invz[khi];
#pragma unroll
                for (int a = 0; a < 4; ++a) sm.imq[a * 128 + e] = rc->mq[a];
                sm.imqd[e] = rc->mqd; sm.ik[e] = rc->k; sm.iorig[e] = rc->orig;
                sm.icm[e] = p.colmax32[ci0 + e];                       // may be one tile stale: still a lower bound of the running maximum
                sm.ipos[e] = ((unsigned long long)rc->pos_hi << 32) | rc->pos_lo;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sfull_bar[buf]);                   // release: the writes above are visible to waiters
            ++it;
        }
    } else {
        // ===================== epilogue warps: four per TMEM lane quadrant, free-running (no CTA barrier) =====================
        const uint32_t q = warp & 3u;                            // TMEM lane quadrant this warp may access
        const uint32_t half = (warp - 2) >> 2;                   // which slice of the 256 accumulator columns
        uint32_t it = 0, ait = 0, r = 0, rph = 0;               // it: record buffers (own valid tiles), ait: accumulator buffers (all tiles)
        I8Emit em; em.cur = I8_INVALID; em.used = 0; em.next = I8_INVALID;
        while (true) {
            q_wait(r, rph);
            const uint2 qt = q_tile[r];
            const uint32_t t0 = __shfl_sync(0xffffffffu, qt.x, 0);                 // consumes the load before the slot is handed back
            if (lane == 0) q_release(r);
            if (++r == I8_QDEPTH) { r = 0; rph ^= 1u; }
            if (t0 == I8_INVALID) break;
            const uint32_t ti = t0 + rank, tj = qt.y;
            const bool valid = own_valid(ti, tj);
            const uint32_t abuf = ait & 1u, acc_phase = (ait >> 1) & 1u;
            const uint32_t buf = it & 1u;
            if (valid) mbar_wait(&sfull_bar[buf], (it >> 1) & 1u);
            mbar_wait(&tfull_bar[abuf], acc_phase);
            tcgen05_fence_after();
#ifdef SPK_I8_SKIP_EPI                                   // experiment: mainloop only (results are wrong)
            if (valid && p.debug_out) {
#else
            if (valid) {
#endif
                I8EpiSmem& sm = *(I8EpiSmem*)(sm_base + buf * SM_STRIDE);
                const uint32_t ci = i8_tile_class(p, ti), cj = i8_tile_class(p, JT * tj);
                const uint32_t WI = 128u >> ci;
                const uint32_t ci0 = p.col_start[ci] + (ti - p.tile_start[ci]) * WI;
                const uint32_t cj0 = p.col_start[cj] + (JT * tj - p.tile_start[cj]) * (128u >> cj);
                const uint32_t tmem_acc = tmem_base + abuf * 256u;
                const bool diag = (ti >= JT * tj);                       // panel touches the diagonal: pairs need pi < pj
                switch (ci * 3u + cj) {
                    case 0: i8_epilogue_tile<1, 1, WR>(p, sm, em, tmem_acc, q, half, lane, diag, ci0, cj0); break;
                    case 1: i8_epilogue_tile<1, 2, WR>(p, sm, em, tmem_acc, q, half, lane, diag, ci0, cj0); break;
                    case 2: i8_epilogue_tile<1, 4, WR>(p, sm, em, tmem_acc, q, half, lane, diag, ci0, cj0); break;
                    case 4: i8_epilogue_tile<2, 2, WR>(p, sm, em, tmem_acc, q, half, lane, diag, ci0, cj0); break;
                    case 5: i8_epilogue_tile<2, 4, WR>(p, sm, em, tmem_acc, q, half, lane, diag, ci0, cj0); break;
                    default: i8_epilogue_tile<4, 4, WR>(p, sm, em, tmem_acc, q, half, lane, diag, ci0, cj0); break;
                }
            }
            // accumulator drained and records consumed: hand both buffers back
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR && !leader) mbar_arrive_remote_nofence(&tempty_bar[abuf], 0); else mbar_arrive(&tempty_bar[abuf]);
                if (valid) mbar_arrive(&sempty_bar[buf]);
            }
            if (valid) ++it;
            ++ait;
        }
        i8_emit_flush(p, em, lane);
    }

    tcgen05_fence_before();
    if constexpr (PAIR) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        if constexpr (PAIR) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// =============================================================================================
// preparation kernels: integer marginals + epilogue tables, operand expansion
// =============================================================================================
// column record from the integer marginals m[5] (by state code; valid in every lane) of permuted column pc: kept-order marginals,
// position, and the entropy / normaliser tables by the partner column's number of present states (lane t computes entry t of the
// five, in fp64; lane 0 writes the record). Called by the whole warp.
__device__ __forceinline__ void i8_write_record(ColRec8* __restrict__ out, uint32_t orig, const uint32_t* m, const uint32_t* __restrict__ info,
                                                const uint64_t* __restrict__ pos, uint32_t pc, double wq_total, double pcq, uint32_t lane) {
    ColRec8 r; memset(&r, 0, sizeof r);
    r.orig = orig;
    uint32_t k = 0;
    double pm[NSTATES] = {0.0, 0.0, 0.0, 0.0, 0.0}; int npres = 0;      // marginals of the present states
    if (orig != I8_INVALID) {
        const uint32_t inf = info[pc];
        k = inf & 15u;
#pragma unroll
        for (uint32_t a = 0; a < 4; ++a)
            if (a < k) {
                const uint32_t code = (inf >> (4 + 3 * a)) & 7u;
                const uint32_t v = code == 0 ? m[0] : code == 1 ? m[1] : code == 2 ? m[2] : code == 3 ? m[3] : m[4];
                r.mq[a] = v; r.smq += v; pm[a] = (double)v;
            }
        const uint32_t drop = (inf >> 16) & 7u;
        r.mqd = drop == 0 ? m[0] : drop == 1 ? m[1] : drop == 2 ? m[2] : drop == 3 ? m[3] : m[4];
        npres = (int)k + 1;
        const uint64_t ps = pos[orig];
        r.pos_lo = (uint32_t)ps; r.pos_hi = (uint32_t)(ps >> 32);
    }
    r.k = k;
    const int mm = (int)(lane % NSTATES) + 1;
    const double Z = wq_total + pcq * (double)(k + 1) * (double)mm;
    double h = 0.0;
#pragma unroll
    for (int a = 0; a < 4; ++a)
        if (a < (int)k) { const double pr = (pm[a] + pcq * (double)mm) / Z; h += pr * log(pr); }
    if (npres) { const double pr = ((double)r.mqd + pcq * (double)mm) / Z; h += pr * log(pr); }
    const float hf = (float)h, izf = (float)(1.0 / Z);
#pragma unroll
    for (int t = 0; t < NSTATES; ++t) { r.h[t] = __shfl_sync(0xffffffffu, hf, t); r.invz[t] = __shfl_sync(0xffffffffu, izf, t); }
    if (lane == 0) *out = r;
}

// one warp per permuted column (records only: used when just the pseudocount changed; i8_prep_kernel is the normal path)
__global__ void i8_colrec_kernel(const uint8_t* __restrict__ colmajor, const uint32_t* __restrict__ qw,
                                 const uint32_t* __restrict__ perm, const uint32_t* __restrict__ info,
                                 const uint64_t* __restrict__ pos, ColRec8* __restrict__ rec,
                                 uint32_t Lperm, uint32_t n, uint32_t n_pad, double wq_total, double pcq) {
    const uint32_t pc = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    const uint32_t lane = threadIdx.x % 32;
    if (pc >= Lperm) return;
    const uint32_t orig = perm[pc];
    uint32_t m[NSTATES] = {0, 0, 0, 0, 0};
    if (orig != I8_INVALID) {
        const uint8_t* col = colmajor + (size_t)orig * n_pad;
        for (uint32_t s = lane; s < n; s += 32) {
            const uint32_t x = col[s], w = qw[s];
#pragma unroll
            for (int a = 0; a < NSTATES; ++a) m[a] += (x == (uint32_t)a) ? w : 0u;
        }
#pragma unroll
        for (int a = 0; a < NSTATES; ++a)
            for (int o = 16; o; o >>= 1) m[a] += __shfl_xor_sync(0xffffffffu, m[a], o);
    }
    i8_write_record(rec + pc, orig, m, info, pos, pc, wq_total, pcq, lane);
}

// Operands AND column records in one pass over the column-major alignment: one warp per permuted column, each lane converts 16
// consecutive samples per iteration. wr = 2: A = one-hot, B = one-hot x {hi, lo} weight byte (two rows per kept state); wr = 1:
// A = one-hot x a_s (byte array qhi), B = one-hot x b_s (byte array qlo), one row per kept state at the same row index as in A.
// EVERY row of the column's class slot is written (zeros for rows the column does not use, for padding columns and for the padded
// samples), so the operand arrays need no memset: 1.1 GB written once per step instead of twice, the 0.3 GB alignment read once.
__global__ void __launch_bounds__(256)
i8_prep_kernel(const uint8_t* __restrict__ colmajor, const uint8_t* __restrict__ qhi, const uint8_t* __restrict__ qlo,
               const uint32_t* __restrict__ perm, const uint32_t* __restrict__ info, const uint64_t* __restrict__ pos,
               uint8_t* __restrict__ A8, uint8_t* __restrict__ B8, ColRec8* __restrict__ rec, uint32_t Lperm, uint32_t n, uint32_t n_pad,
               uint32_t cs1, uint32_t cs2, uint32_t ts1, uint32_t ts2, uint32_t wr, double wq_total, double pcq) {
    const uint32_t pc = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (pc >= Lperm) return;
    const uint32_t orig = perm[pc];
    const uint32_t lane = threadIdx.x % 32;
    const uint32_t cls = (pc >= cs2) ? 2u : (pc >= cs1) ? 1u : 0u;
    const uint32_t kp = 1u << cls;
    const uint32_t cstart = cls == 2 ? cs2 : cls == 1 ? cs1 : 0u;
    const uint32_t tstart = cls == 2 ? ts2 : cls == 1 ? ts1 : 0u;
    const uint32_t lc = pc - cstart;
    const size_t rowA = (size_t)128 * tstart + (size_t)lc * kp;
    const size_t rowB = (wr == 2) ? (size_t)256 * tstart + (size_t)lc * 2 * kp : rowA;
    const bool live = orig != I8_INVALID;
    const uint32_t inf = live ? info[pc] : 0u;
    const uint32_t k = inf & 15u;
    const uint8_t* col = colmajor + (size_t)(live ? orig : 0u) * n_pad;
    // integer marginals of the kept states come from the operand bytes just built (a_s * b_s resp. hi * 256 + lo of the samples
    // in that state, 0 elsewhere); the dropped state's is the total minus their sum
    uint32_t mk[4] = {0, 0, 0, 0};
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    // byte-wise SIMD on 32-bit words (__vcmpeq4 / AND / __dp4a): samples beyond n carry weight bytes 0 (and the GAP code), so they
    // contribute nothing and need no bound test
    for (uint32_t s0 = lane * 16; s0 < n_pad; s0 += 512) {
        uint4 raw = zero, rh = zero, rl = zero;
        if (live) {
            raw = *reinterpret_cast<const uint4*>(col + s0);
            rh = *reinterpret_cast<const uint4*>(qhi + s0);
            rl = *reinterpret_cast<const uint4*>(qlo + s0);
        }
        const uint32_t xw[4] = {raw.x, raw.y, raw.z, raw.w}, hw[4] = {rh.x, rh.y, rh.z, rh.w}, lw[4] = {rl.x, rl.y, rl.z, rl.w};
#pragma unroll
        for (uint32_t a = 0; a < 4; ++a) {
            if (a >= kp) break;
            uint32_t oa[4] = {0, 0, 0, 0}, oh[4] = {0, 0, 0, 0}, ol[4] = {0, 0, 0, 0};
            if (live && a < k) {
                const uint32_t code4 = ((inf >> (4 + 3 * a)) & 7u) * 0x01010101u;
                uint32_t acc = 0, acc_hi = 0;
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const uint32_t hit = __vcmpeq4(xw[w], code4);                 // 0xff in the bytes of samples in this state
                    oh[w] = hit & hw[w]; ol[w] = hit & lw[w];
                    oa[w] = (wr == 2) ? (hit & 0x01010101u) : oh[w];
                    if (wr == 2) { acc_hi = __dp4a(oh[w], 0x01010101u, acc_hi); acc = __dp4a(ol[w], 0x01010101u, acc); }
                    else acc = __dp4a(oh[w], ol[w], acc);
                }
                mk[a] += acc + (acc_hi << 8);
            }
            *reinterpret_cast<uint4*>(A8 + (rowA + a) * n_pad + s0) = make_uint4(oa[0], oa[1], oa[2], oa[3]);
            if (wr == 2) {
                *reinterpret_cast<uint4*>(B8 + (rowB + a) * n_pad + s0) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
                *reinterpret_cast<uint4*>(B8 + (rowB + kp + a) * n_pad + s0) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
            } else {
                *reinterpret_cast<uint4*>(B8 + (rowB + a) * n_pad + s0) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a)
        for (int o = 16; o; o >>= 1) mk[a] += __shfl_xor_sync(0xffffffffu, mk[a], o);
    uint32_t m[NSTATES] = {0, 0, 0, 0, 0};
    if (live) {
        uint32_t sum = 0;
#pragma unroll
        for (uint32_t a = 0; a < 4; ++a)
            if (a < k) {
                const uint32_t code = (inf >> (4 + 3 * a)) & 7u;
#pragma unroll
                for (int t = 0; t < NSTATES; ++t) if (code == (uint32_t)t) m[t] = mk[a];
                sum += mk[a];
            }
        const uint32_t drop = (inf >> 16) & 7u, md = (uint32_t)((unsigned long long)wq_total - sum);
#pragma unroll
        for (int t = 0; t < NSTATES; ++t) if (drop == (uint32_t)t) m[t] = md;
    }
    i8_write_record(rec + pc, orig, m, info, pos, pc, wq_total, pcq, lane);
}

__global__ void i8_fill_u32_kernel(uint32_t* p, uint32_t v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// =============================================================================================
// host side
// =============================================================================================
template <class P> static void i8_free_dev(P*& p) { if (p) { cudaFree(p); p = nullptr; } }

void i8_free(Ctx* c) {
    i8_free_dev(c->d_i8_perm); i8_free_dev(c->d_i8_info); i8_free_dev(c->d_i8_rec); i8_free_dev(c->d_A8); i8_free_dev(c->d_B8);
    i8_free_dev(c->d_q); i8_free_dev(c->d_i8_colmax32); i8_free_dev(c->d_i8_colmax32_seed); c->i8_seed_valid = false;
    c->i8 = I8Layout();
    c->i8_weights_valid = c->i8_operands_valid = false; c->i8_rec_pc = -1.0;
    c->i8_cap_Lperm = c->i8_cap_NT = c->i8_cap_npad = 0;
}

// classes + permutation from the presence masks (host; O(L))
int i8_build_layout(Ctx* c) {
    const uint32_t L = c->L;
    c->h_masks.resize(L);
    SPK_CUDA(c, cudaMemcpyAsync(c->h_masks.data(), c->d_masks, L, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    I8Layout lay;
    std::vector<uint8_t> cls(L);
    for (uint32_t col = 0; col < L; ++col) {
        const int np = __builtin_popcount(c->h_masks[col] & 31u);
        const int k = np > 0 ? np - 1 : 0;
        cls[col] = (uint8_t)(k <= 1 ? 0 : (k == 2 ? 1 : 2));
        lay.ncols[cls[col]]++;
    }
    uint32_t cstart = 0, tstart = 0;
    for (int g = 0; g < 3; ++g) {
        const uint32_t width = 128u >> g;
        lay.ntiles[g] = (lay.ncols[g] + width - 1) / width;
        lay.ntiles[g] += lay.ntiles[g] & 1u;                     // even: a 256-row B panel (two column tiles) never straddles classes
        lay.ncols_pad[g] = lay.ntiles[g] * width;
        lay.col_start[g] = cstart; lay.tile_start[g] = tstart;
        cstart += lay.ncols_pad[g]; tstart += lay.ntiles[g];
    }
    lay.Lperm = cstart; lay.NT = tstart;
    std::vector<uint32_t> perm(lay.Lperm, I8_INVALID), info(lay.Lperm, 0u);
    uint32_t fill[3] = {lay.col_start[0], lay.col_start[1], lay.col_start[2]};
    for (uint32_t col = 0; col < L; ++col) {
        const uint32_t pc = fill[cls[col]]++;
        perm[pc] = col;
        const uint32_t mask = c->h_masks[col] & 31u;
        int drop = 4; while (drop > 0 && !((mask >> drop) & 1u)) --drop;          // highest-coded present state
        uint32_t inf = 0, k = 0;
        for (int a = 0; a < drop; ++a) if ((mask >> a) & 1u) { inf |= (uint32_t)a << (4 + 3 * k); ++k; }
        info[pc] = inf | k | ((uint32_t)drop << 16);
    }
    // re-uploads of the same shape keep their device buffers (cudaFree/cudaMalloc of GBs costs tens of ms)
    if (lay.Lperm > c->i8_cap_Lperm || lay.NT > c->i8_cap_NT || c->n_pad != c->i8_cap_npad) i8_free(c);
    c->i8 = lay;
    c->i8_weights_valid = c->i8_operands_valid = false; c->i8_rec_pc = -1.0;
    if (c->i8_cap_Lperm == 0) {
    c->i8_cap_Lperm = lay.Lperm + 1024; c->i8_cap_NT = lay.NT + 8; c->i8_cap_npad = c->n_pad;
    const uint32_t capL = c->i8_cap_Lperm;
    if (cudaMalloc(&c->d_i8_perm, (size_t)capL * 4) != cudaSuccess || cudaMalloc(&c->d_i8_info, (size_t)capL * 4) != cudaSuccess ||
        cudaMalloc(&c->d_i8_rec, (size_t)capL * sizeof(ColRec8)) != cudaSuccess ||
        cudaMalloc(&c->d_i8_colmax32, (size_t)capL * 4) != cudaSuccess || cudaMalloc(&c->d_i8_colmax32_seed, (size_t)capL * 4) != cudaSuccess ||
        cudaMalloc(&c->d_q, (size_t)c->n_pad * 6) != cudaSuccess) {
        cudaGetLastError();
        return fail(c, SPK_E_NOMEM, "int8 engine: layout allocation failed");
    }
    }
    SPK_CUDA(c, cudaMemcpyAsync(c->d_i8_perm, perm.data(), (size_t)lay.Lperm * 4, cudaMemcpyHostToDevice, c->stream));
    SPK_CUDA(c, cudaMemcpyAsync(c->d_i8_info, info.data(), (size_t)lay.Lperm * 4, cudaMemcpyHostToDevice, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}

// nearest product a*b (1 <= a, b <= 255) to every target 0..65025: table built once
static const std::vector<uint16_t>& i8_product_table() {
    static std::vector<uint16_t> tab;                    // tab[t] = (a << 8) | b with a*b nearest to t (0 for t = 0)
    if (tab.empty()) {
        const int T = 255 * 255;
        std::vector<int> best(T + 1, -1);
        for (int a = 1; a <= 255; ++a) for (int b = a; b <= 255; ++b) best[a * b] = (a << 8) | b;
        std::vector<uint16_t> t2(T + 1, 0);
        int last = -1;
        std::vector<int> prev(T + 1, -1), next(T + 1, -1);
        for (int t = 0; t <= T; ++t) { if (best[t] >= 0) last = t; prev[t] = last; }
        last = -1;
        for (int t = T; t >= 0; --t) { if (best[t] >= 0) last = t; next[t] = last; }
        for (int t = 1; t <= T; ++t) {
            const int lo = prev[t], hi = next[t];
            const int pick = (lo < 1) ? hi : (hi < 0 ? lo : ((t - lo) <= (hi - t) ? lo : hi));
            t2[t] = (uint16_t)best[pick];
        }
        tab.swap(t2);
    }
    return tab;
}

// weight quantisation (host; O(n)), operands, column records.
// wr = 2: q_s = round(w_s / w_max * 65535), the N operand carries the two bytes of q_s (exact 16-bit grid);
// wr = 1: q_s = a_s * b_s, the nearest product of two bytes to w_s / w_max * M (M chosen among a few scales for the
//         smallest total rounding error): the M operand carries a_s, the N operand b_s -- half the N rows, i.e.
//         half the tensor work per column pair, on a slightly coarser (but still exact-integer) weight grid.
static int i8_prepare(Ctx* c, double pc, int wr) {
    const uint32_t n = c->n, n_pad = c->n_pad;
    if (n > 65536u) return fail(c, SPK_E_INVALID, "int8 engine supports at most 65536 samples (s32 accumulators); use engine 2");
    if (c->i8.NT == 0) return fail(c, SPK_E_INVALID, "int8 engine: no layout (set_alignment first)");
    uint8_t* d_qhi = (uint8_t*)(c->d_q + n_pad); uint8_t* d_qlo = d_qhi + n_pad;
    if (c->i8_wr != wr) { c->i8_weights_valid = false; c->i8_wr = wr; }
    if (!c->i8_weights_valid) {
        if (c->h_w.size() != n) return fail(c, SPK_E_INVALID, "int8 engine: weights not set");
        double wmax = 0.0;
        for (uint32_t s = 0; s < n; ++s) {
            if (!(c->h_w[s] >= 0.0) || !std::isfinite(c->h_w[s])) return fail(c, SPK_E_INVALID, "int8 engine: weights must be finite and >= 0");
            wmax = std::max(wmax, c->h_w[s]);
        }
        if (!(wmax > 0.0)) return fail(c, SPK_E_INVALID, "int8 engine: all weights are zero");
        std::vector<uint32_t> q(n_pad, 0u); std::vector<uint8_t> qb(2 * (size_t)n_pad, 0);
        unsigned long long tot = 0;
        if (wr == 2) {
            for (uint32_t s = 0; s < n; ++s) {
                const uint32_t v = (uint32_t)std::llround(c->h_w[s] / wmax * 65535.0);
                q[s] = v; qb[s] = (uint8_t)(v >> 8); qb[n_pad + s] = (uint8_t)(v & 255u); tot += v;
            }
            c->i8_delta = wmax / 65535.0;
        } else {
            const std::vector<uint16_t>& tab = i8_product_table();
            // sample weights take few distinct values (1 / #neighbours): the scale search runs over (value, multiplicity)
            std::vector<double> sorted_w(c->h_w.begin(), c->h_w.begin() + n);
            std::sort(sorted_w.begin(), sorted_w.end());
            std::vector<std::pair<double, double>> distinct;                      // (w / w_max, count)
            for (uint32_t s = 0; s < n;) {
                uint32_t e = s;
                while (e < n && sorted_w[e] == sorted_w[s]) ++e;
                distinct.emplace_back(sorted_w[s] / wmax, (double)(e - s));
                s = e;
            }
            double best_err = 1e300, best_M = 65025.0;
            for (int trial = 0; trial < 24; ++trial) {
                const double M = 65025.0 * (1.0 - 0.02 * trial);                  // scales between 52 % and 100 % of the byte-product range
                double err = 0.0;
                for (const auto& d : distinct) {
                    const double t = d.first * M;
                    const long ti = std::min<long>(65025, std::llround(t));
                    const uint16_t ab = tab[ti];
                    const double qv = (double)(ab >> 8) * (double)(ab & 255u);
                    err += d.second * (std::fabs(qv - t) / M);                     // in units of w_max
                }
                if (err < best_err) { best_err = err; best_M = M; }
            }
            for (uint32_t s = 0; s < n; ++s) {
                const long ti = std::min<long>(65025, std::llround(c->h_w[s] / wmax * best_M));
                const uint16_t ab = tab[ti];
                const uint32_t a = ab >> 8, b = ab & 255u;
                q[s] = a * b; qb[s] = (uint8_t)a; qb[n_pad + s] = (uint8_t)b; tot += q[s];
            }
            c->i8_delta = wmax / best_M;
        }
        c->i8_wq_total = tot;
        // (pageable sources: cudaMemcpyAsync stages them before it returns, so the vectors may go out of scope)
        SPK_CUDA(c, cudaMemcpyAsync(c->d_q, q.data(), (size_t)n_pad * 4, cudaMemcpyHostToDevice, c->stream));
        SPK_CUDA(c, cudaMemcpyAsync(d_qhi, qb.data(), (size_t)n_pad * 2, cudaMemcpyHostToDevice, c->stream));
        c->i8_weights_valid = true; c->i8_operands_valid = false; c->i8_rec_pc = -1.0;
    }
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_EXP0], c->stream));
    if (!c->i8_operands_valid) {
        const size_t rowsA = (size_t)128 * c->i8.NT;
        if (!c->d_A8) {
            const size_t capA = (size_t)128 * c->i8_cap_NT;
            if (cudaMalloc(&c->d_A8, capA * n_pad) != cudaSuccess || cudaMalloc(&c->d_B8, 2 * capA * n_pad) != cudaSuccess) {
                cudaGetLastError(); i8_free_dev(c->d_A8);
                return fail(c, SPK_E_NOMEM, "int8 engine: cannot allocate operands (%zu bytes)", 3 * rowsA * n_pad);
            }
        }
        // (no memset: i8_prep_kernel writes every row of every class slot, padding included)
        i8_prep_kernel<<<(c->i8.Lperm + 7) / 8, 256, 0, c->stream>>>(c->d_colmajor, d_qhi, d_qlo, c->d_i8_perm, c->d_i8_info, c->d_pos, c->d_A8, c->d_B8, c->d_i8_rec,
                                                                    c->i8.Lperm, n, n_pad, c->i8.col_start[1], c->i8.col_start[2],
                                                                    c->i8.tile_start[1], c->i8.tile_start[2], (uint32_t)wr,
                                                                    (double)c->i8_wq_total, pc / c->i8_delta);
        c->tm.kernel_launches += 1;
        SPK_CUDA(c, cudaGetLastError());
        c->i8_rec_pc = pc;
        c->i8_operands_valid = true;
    }
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_EXP1], c->stream));
    if (c->i8_rec_pc != pc) {
        i8_colrec_kernel<<<(c->i8.Lperm + 7) / 8, 256, 0, c->stream>>>(c->d_colmajor, c->d_q, c->d_i8_perm, c->d_i8_info, c->d_pos, c->d_i8_rec,
                                                                      c->i8.Lperm, n, n_pad, (double)c->i8_wq_total, pc / c->i8_delta);
        c->tm.kernel_launches += 1;
        SPK_CUDA(c, cudaGetLastError());
        c->i8_rec_pc = pc;
    }
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_STAT1], c->stream));
    c->tm_pending |= TMP_EXPAND | TMP_STATS;
    return SPK_OK;
}

typedef CUresult (*PFN_encodeTiled8)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 2D u8 tensor [rows][n_pad] (samples contiguous), box = 128 samples x box_rows, 128B swizzle
static int i8_make_map(Ctx* c, CUtensorMap* map, void* base, uint64_t rows, uint64_t k_bytes, uint32_t box_rows) {
    static PFN_encodeTiled8 enc = nullptr;
    if (!enc) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            enc = (PFN_encodeTiled8)ptr;
    }
    if (!enc) return fail(c, SPK_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t gdim[2] = {k_bytes, rows};
    cuuint64_t gstride[1] = {k_bytes};
    cuuint32_t box[2] = {I8_BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(c, SPK_E_CUDA, "cuTensorMapEncodeTiled (u8) failed (%d)", (int)r);
    return SPK_OK;
}

template <int NCTA, int WR> static size_t i8_smem_bytes() {
    return (size_t)I8Cfg<NCTA, WR>::NSTAGES * I8Cfg<NCTA, WR>::STAGE + 2 * ((sizeof(I8EpiSmemT<256 / WR>) + 15) & ~(size_t)15) +
           (2 * I8Cfg<NCTA, WR>::NSTAGES + 8 + 3 * I8_QDEPTH) * 8 + 16;
}

// SPYDRPICK_B200_I8_PAIR=0 selects the one-CTA-per-tile kernel (A/B comparisons); default: CTA pairs
static bool i8_use_pair() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("SPYDRPICK_B200_I8_PAIR"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

// SPYDRPICK_B200_I8_SEED=0 disables the colmax seed pass (A/B comparisons)
constexpr uint32_t I8_SEED_STRIDE = 32;
#ifndef SPK_I8_DEFAULT_WR
#define SPK_I8_DEFAULT_WR 1
#endif
constexpr int I8_DEFAULT_WR = SPK_I8_DEFAULT_WR;
static bool i8_use_seed() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("SPYDRPICK_B200_I8_SEED"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

template <int WR> static int i8_launch_wr(Ctx* c, I8Params& gp) {
    const bool pair = gp.pair != 0;
    CUtensorMap ma, mb;
    int r = i8_make_map(c, &ma, c->d_A8, (uint64_t)128 * c->i8.NT, c->n_pad, 128); if (r) return r;
    r = i8_make_map(c, &mb, c->d_B8, (uint64_t)(128 * WR) * c->i8.NT, c->n_pad, pair ? 128 : 256); if (r) return r;
    gp.rec = c->d_i8_rec;
    gp.num_kb = c->n_pad / I8_BK;
    gp.NT = c->i8.NT;
    gp.jt = 2u / WR; gp.NTJ = c->i8.NT / gp.jt;
    for (int g = 0; g < 3; ++g) { gp.tile_start[g] = c->i8.tile_start[g]; gp.col_start[g] = c->i8.col_start[g]; }
    gp.colmax32 = c->d_i8_colmax32;
    gp.counters = c->d_counters;
    if (!gp.cand_counter) { gp.cand_counter = c->d_counters + 0; gp.cand_overflow = c->d_counters + 4; }
    const unsigned long long slots = gp.debug_out ? 1ull : gp.n_local_slots;
    if (!pair) {
        const unsigned grid = (unsigned)(slots < (unsigned long long)c->sm_count ? (slots ? slots : 1) : (unsigned long long)c->sm_count);
        const size_t smem = i8_smem_bytes<1, WR>();
        SPK_CUDA(c, cudaFuncSetAttribute(mi_i8_kernel<1, WR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        mi_i8_kernel<1, WR><<<grid, I8_THREADS, smem, c->stream>>>(ma, mb, gp);
    } else {
        const size_t smem = i8_smem_bytes<2, WR>();
        SPK_CUDA(c, cudaFuncSetAttribute(mi_i8_kernel<2, WR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof cfg);
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.blockDim = dim3(I8_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream; cfg.attrs = attr; cfg.numAttrs = 1;
        cfg.gridDim = dim3(2u * (unsigned)(c->sm_count / 2));
        int max_clusters = 0;
        if (cudaOccupancyMaxActiveClusters(&max_clusters, mi_i8_kernel<2, WR>, &cfg) != cudaSuccess || max_clusters < 1) {
            cudaGetLastError();
            max_clusters = c->sm_count / 2;
        }
        unsigned long long pairs = (unsigned long long)max_clusters;
        if (slots < pairs) pairs = slots ? slots : 1;
        cfg.gridDim = dim3(2u * (unsigned)pairs);
        SPK_CUDA(c, cudaLaunchKernelEx(&cfg, mi_i8_kernel<2, WR>, ma, mb, gp));
    }
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}
static int i8_launch(Ctx* c, I8Params& gp, int wr) { return wr == 2 ? i8_launch_wr<2>(c, gp) : i8_launch_wr<1>(c, gp); }

// weight rows of the default int8 engine (engine 0): SPYDRPICK_B200_I8_WR=2 selects the exact 16-bit {hi, lo} rows,
// =1 the factorised single-row weights
int i8_default_wr() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("SPYDRPICK_B200_I8_WR"); v = (e && e[0] == '2') ? 2 : ((e && e[0] == '1') ? 1 : I8_DEFAULT_WR); }
    return v;
}

int mi_i8_filter(Ctx* c, const spk_scan_params& p, double band, uint64_t circle, int wr, bool reuse_seed) {
    int r = i8_prepare(c, p.pseudocount, wr); if (r) return r;
    I8Params gp; memset(&gp, 0, sizeof gp);
    const uint32_t NT = c->i8.NT;
    const uint32_t JT = 2u / (uint32_t)wr, NTJ = NT / JT, SJ = I8_SUPER / JT;      // B panels; panels per super-row step
    const uint32_t n_sr = (NT + I8_SUPER - 1) / I8_SUPER;
    std::vector<unsigned long long> prefix(n_sr + 1, 0);
    for (uint32_t sr = 0; sr < n_sr; ++sr) prefix[sr + 1] = prefix[sr] + (NTJ - sr * SJ);
    if ((size_t)(n_sr + 1) * 8 > (1u << 20)) return fail(c, SPK_E_INVALID, "int8 engine: too many column tiles");
    const uint32_t nparts = p.n_parts ? p.n_parts : 1;
    if (p.part_index >= nparts) return fail(c, SPK_E_INVALID, "part_index >= n_parts");
    const unsigned long long total_chunks = prefix[n_sr];
    const unsigned long long local_chunks = total_chunks > p.part_index ? (total_chunks - p.part_index + nparts - 1) / nparts : 0;
    unsigned long long* d_prefix = (unsigned long long*)c->d_tilemap;
    SPK_CUDA(c, cudaMemcpyAsync(d_prefix, prefix.data(), (n_sr + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    gp.cand = c->d_cand; gp.cand_cap = c->cand_cap;
    gp.sr_prefix = d_prefix; gp.n_sr = n_sr;
    gp.pair = i8_use_pair() ? 1u : 0u;
    gp.part = p.part_index; gp.n_parts = nparts; gp.n_local_slots = local_chunks * (gp.pair ? I8_SUPER / 2 : I8_SUPER);
    gp.pcq = (float)(p.pseudocount / c->i8_delta);
    gp.band = (float)band; gp.thr_lo = (float)(p.threshold - band); gp.thr_hi = (float)(p.threshold + band);
    gp.ld_threshold = p.ld_threshold; gp.circle = circle; gp.linear = p.linear_genome;
    gp.debug_out = nullptr;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_GEMM0], c->stream));
    // Seed pass: a 1/I8_SEED_STRIDE sample of the chunks (of ALL partitions), to bring every column's running maximum close to
    // its final value before the main pass. Without it the open-column rule (candidates within `band` of a column's running
    // maximum while that maximum is below thr + band) fires on every record of a rising maximum: measured 6.6 M extra
    // candidates and 5.5 ms of slow-path epilogue per launch at 3000 x 100000. The same launch emits the AUDIT sample of the
    // filter (pairs with mi32 > thr - 4*band of the sampled chunks) into its own buffer.
    // Multi-GPU (c->world > 1 and this scan is the collective one): the sampled chunks are dealt to the ranks and the maxima
    // max-reduced over NVLink, instead of every rank repeating the whole seed pass.
    // A rescan of the same scan (wider band, larger buffers) reuses the seeded maxima: it is local to the rank that needs it.
    bool seeded = false;
    const bool collective = c->world > 1 && nparts == (uint32_t)c->world && p.part_index == (uint32_t)c->rank;
    if (reuse_seed && c->i8_seed_valid) {
        SPK_CUDA(c, cudaMemcpyAsync(c->d_i8_colmax32, c->d_i8_colmax32_seed, (size_t)c->i8.Lperm * 4, cudaMemcpyDeviceToDevice, c->stream));
        SPK_CUDA(c, cudaEventRecord(c->ev[EV_SEED1], c->stream));
    } else {
        i8_fill_u32_kernel<<<(c->i8.Lperm + 255) / 256, 256, 0, c->stream>>>(c->d_i8_colmax32, f32_ordered(-INFINITY), c->i8.Lperm);
        c->tm.kernel_launches += 1;
        c->i8_seed_valid = false;
        // small problems: a denser sample (down to every chunk) so that the audit still sees a meaningful number of pairs
        const uint32_t stride = (uint32_t)std::max<unsigned long long>(1, std::min<unsigned long long>(I8_SEED_STRIDE, total_chunks / 64));
        if (i8_use_seed() && total_chunks >= 1) {
            std::vector<unsigned long long> seed;
            unsigned long long k = 0;
            for (uint32_t sr = 0; sr < n_sr; ++sr)
                for (unsigned long long q = prefix[sr]; q < prefix[sr + 1]; ++q)
                    if ((uint32_t)(q - prefix[sr]) % stride == sr % stride) {
                        if (!collective || k % (unsigned long long)c->world == (unsigned long long)c->rank) seed.push_back(q);
                        ++k;
                    }
            if ((size_t)(n_sr + 1 + seed.size()) * 8 <= (1u << 20)) {
                if (!seed.empty()) {
                    SPK_CUDA(c, cudaMemcpyAsync(d_prefix + n_sr + 1, seed.data(), seed.size() * 8, cudaMemcpyHostToDevice, c->stream));
                    I8Params sp = gp;
                    sp.seed_chunks = d_prefix + n_sr + 1;
                    sp.n_local_slots = (unsigned long long)seed.size() * (gp.pair ? I8_SUPER / 2 : I8_SUPER);
                    if (c->d_audit && c->audit_cap >= 4096) {
                        sp.cand = c->d_audit; sp.cand_cap = c->audit_cap;
                        sp.cand_counter = c->d_counters + 11; sp.cand_overflow = c->d_counters + 12;
                        sp.thr_lo = (float)(p.threshold - 4.0 * band);
                    } else {
                        sp.thr_lo = INFINITY;
                    }
                    r = i8_launch(c, sp, wr); if (r) return r;
                    SPK_CUDA(c, cudaMemsetAsync(c->d_counters + 5, 0, sizeof(unsigned long long), c->stream));   // rewind the tile cursor
                }
                if (collective) { r = comm_allreduce_max_u32(c, c->d_i8_colmax32, c->i8.Lperm); if (r) return r; }
                seeded = true;
                if (c->d_i8_colmax32_seed) {
                    SPK_CUDA(c, cudaMemcpyAsync(c->d_i8_colmax32_seed, c->d_i8_colmax32, (size_t)c->i8.Lperm * 4, cudaMemcpyDeviceToDevice, c->stream));
                    c->i8_seed_valid = true;
                }
            }
        }
        SPK_CUDA(c, cudaEventRecord(c->ev[EV_SEED1], c->stream));
    }
    (void)seeded;
    r = i8_launch(c, gp, wr); if (r) return r;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_GEMM1], c->stream));
    c->tm_pending |= TMP_GEMM | TMP_SEED;
    // tiles = M128 x N256 MMA tiles of the main pass; pairs = column pairs they cover (incl. padding / below-diagonal halves)
    unsigned long long tiles = 0, pairs = 0;
    auto width = [&](uint32_t t) { return t >= c->i8.tile_start[2] ? 32u : (t >= c->i8.tile_start[1] ? 64u : 128u); };
    for (unsigned long long lq = 0; lq < local_chunks; ++lq) {
        const unsigned long long q = lq * nparts + p.part_index;
        const uint32_t sr = (uint32_t)(std::upper_bound(prefix.begin(), prefix.end(), q) - prefix.begin()) - 1;
        const uint32_t tj = sr * SJ + (uint32_t)(q - prefix[sr]);
        for (uint32_t w = 0; w < I8_SUPER; ++w) {
            const uint32_t ti = sr * I8_SUPER + w;
            if (ti < NT && tj < NTJ && ti <= JT * tj + (JT - 1)) { ++tiles; pairs += (unsigned long long)width(ti) * JT * width(JT * tj); }
        }
    }
    c->tm.gemm_tiles = tiles; c->tm.gemm_pairs = pairs;
    return SPK_OK;
}

// raw integer tables of one tile pair: out[(128/KI)][(128/KJ)][KI][KJ] (<= 16384 words)
int mi_i8_debug_tile(Ctx* c, uint32_t ti, uint32_t tj, double pc, uint32_t* d_out) {
    const int wr = i8_default_wr();
    int r = i8_prepare(c, pc, wr); if (r) return r;
    if (ti > tj || tj >= c->i8.NT) return fail(c, SPK_E_INVALID, "int8 debug tile: need ti <= tj < NT");
    I8Params gp; memset(&gp, 0, sizeof gp);
    gp.sr_prefix = (const unsigned long long*)c->d_tilemap;
    gp.pair = i8_use_pair() ? 1u : 0u;
    gp.n_parts = 1; gp.n_local_slots = 1;
    gp.pcq = (float)(pc / c->i8_delta);
    gp.debug_out = d_out; gp.debug_ti = ti; gp.debug_tj = tj;
    r = i8_launch(c, gp, wr); if (r) return r;
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}


// =============================================================================================
// Tensor-pipe peak of the instruction the filter issues (the roofline denominator bench.py reports against).
// A CTA pair per TPC issues tcgen05.mma.cta_group::2.kind::i8 M256 N256 K32 back to back on operands that already sit in
// shared memory (same 128B-swizzled K-major stages, same 4 MMAs per 128-byte K block, two alternating accumulators): no TMA,
// no epilogue, no L2 traffic -- what the pipe sustains when nothing else limits it. cuBLASLt's s8 GEMM (torch._int_mm) reaches
// only ~75 % of this figure, which is why the filter's issued rate exceeded that "peak".
// =============================================================================================
constexpr int I8_PEAK_STAGES = 5;
constexpr uint32_t I8_PEAK_STAGE_BYTES = I8_A_BYTES + 128u * I8_BK;        // 16 KB A + this CTA's half of B
constexpr uint32_t I8_PEAK_BATCH = 64;                                      // K blocks per commit

__global__ void __launch_bounds__(128, 1)
i8_peak_kernel(uint32_t kblocks) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t batch_bar[2];
    __shared__ __align__(8) uint64_t fin_bar;
    __shared__ uint32_t tmem_slot;
    const uint32_t warp = threadIdx.x / 32, rank = cluster_ctarank();
    // operand bytes: any pattern (the accumulators wrap, nothing is read back); small values as the one-hot operands have
    for (uint32_t e = threadIdx.x; e < I8_PEAK_STAGES * I8_PEAK_STAGE_BYTES / 4; e += blockDim.x)
        ((uint32_t*)smem_raw)[e] = ((e * 2654435761u) >> 7) & 0x01030107u;
    if (threadIdx.x == 0) {
        mbar_init(&batch_bar[0], 1); mbar_init(&batch_bar[1], 1); mbar_init(&fin_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");              // generic-proxy stores -> visible to the tensor core's async proxy
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    cluster_sync_all();
    tcgen05_fence_after();
    const uint32_t tmem_base = tmem_slot;
    if (warp == 1 && rank == 0) {
        constexpr uint32_t IDESC = (I8_IDESC & ~(31u << 24)) | ((256u >> 4) << 24);
        const uint32_t nb = (kblocks + I8_PEAK_BATCH - 1) / I8_PEAK_BATCH;
        uint32_t kb = 0;
        for (uint32_t b = 0; b < nb; ++b) {
            if (b >= 2) mbar_wait(&batch_bar[b & 1u], ((b - 2u) >> 1) & 1u);     // at most two batches in flight
            tcgen05_fence_after();
            if (elect_one()) {
                for (uint32_t e = 0; e < I8_PEAK_BATCH && kb + e < kblocks; ++e) {
                    const uint32_t stage = (kb + e) % I8_PEAK_STAGES;
                    const uint32_t sb = smem_u32(smem_raw + (size_t)stage * I8_PEAK_STAGE_BYTES);
                    const uint64_t db = umma_smem_desc(sb), da = umma_smem_desc(sb + 128u * I8_BK);
                    const uint32_t d_tmem = tmem_base + (((kb + e) / 24u) & 1u) * 256u;      // 24 K blocks = one 3,000-sample tile
#pragma unroll
                    for (uint32_t k = 0; k < I8_BK / 32; ++k) umma_i8_pair(d_tmem, da + 2u * k, db + 2u * k, IDESC, 1u);
                }
                umma_commit_pair(&batch_bar[b & 1u]);
            }
            __syncwarp();
            kb += I8_PEAK_BATCH;
        }
        if (elect_one()) umma_commit_pair(&fin_bar);
        __syncwarp();
    }
    if (warp == 1) { mbar_wait_cluster(&fin_bar, 0u); tcgen05_fence_after(); }
    tcgen05_fence_before();
    cluster_sync_all();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// launches the peak kernel `reps` times back to back; *ms = total device time, *tops = tensor ops per second over that interval
int i8_tensor_peak(Ctx* c, uint32_t kblocks, uint32_t reps, float* ms, double* tops) {
    if (kblocks == 0 || reps == 0) return fail(c, SPK_E_INVALID, "i8_tensor_peak: kblocks and reps must be > 0");
    const size_t smem = (size_t)I8_PEAK_STAGES * I8_PEAK_STAGE_BYTES;
    SPK_CUDA(c, cudaFuncSetAttribute(i8_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg; memset(&cfg, 0, sizeof cfg);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = c->stream; cfg.attrs = attr; cfg.numAttrs = 1;
    cfg.gridDim = dim3(2u * (unsigned)(c->sm_count / 2));
    int max_clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&max_clusters, i8_peak_kernel, &cfg) != cudaSuccess || max_clusters < 1) { cudaGetLastError(); max_clusters = c->sm_count / 2; }
    const unsigned pairs = (unsigned)std::min(max_clusters, c->sm_count / 2);
    cfg.gridDim = dim3(2u * pairs);
    cudaEvent_t e0, e1;
    SPK_CUDA(c, cudaEventCreate(&e0)); SPK_CUDA(c, cudaEventCreate(&e1));
    SPK_CUDA(c, cudaEventRecord(e0, c->stream));
    for (uint32_t r = 0; r < reps; ++r) SPK_CUDA(c, cudaLaunchKernelEx(&cfg, i8_peak_kernel, kblocks));
    SPK_CUDA(c, cudaEventRecord(e1, c->stream));
    SPK_CUDA(c, cudaEventSynchronize(e1));
    float t = 0.f;
    SPK_CUDA(c, cudaEventElapsedTime(&t, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    c->tm.kernel_launches += reps;
    if (ms) *ms = t;
    if (tops) *tops = 2.0 * 256.0 * 256.0 * 32.0 * (I8_BK / 32) * (double)kblocks * pairs * reps / (t * 1e-3) / 1e12;
    return SPK_OK;
}

} // namespace spk
