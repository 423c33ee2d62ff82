// spk_mi_exact.cu -- fp64 exact MI kernels: K2b (pair list), K2r/K2c (refine of filter
// candidates incl. the literal "MI without gaps" second pass), threshold/colmax compaction and
// the exhaustive diagnostic engine.
//
// Formula: include/mi.hpp:146-181 (normalize_and_get_mi_single) on the 5x5 table
// buffer[jstate*5+istate] (misc/Crosstable.hpp:62-67). One warp per column pair: lanes stride the
// samples and accumulate into lane-private shared-memory bins (no atomics, deterministic), then
// lane l < 25 owns cell l = b*5+a and the entropy terms are evaluated one log per lane.
#include <cfloat>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include "spk_internal.h"

namespace spk {

constexpr int EX_WARPS = 4;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// weighted crosstable of two column-major byte columns; returns cell value for lane < 25.
__device__ __forceinline__ double warp_crosstab(const uint8_t* __restrict__ ci, const uint8_t* __restrict__ cj,
                                                const double* __restrict__ w, uint32_t n_pad, double* bins, uint32_t lane) {
#pragma unroll
    for (int k = 0; k < 25; ++k) bins[k * 32 + lane] = 0.0;
    for (uint32_t s0 = lane * 4; s0 < n_pad; s0 += 128) {
        const uchar4 xa = *reinterpret_cast<const uchar4*>(ci + s0);
        const uchar4 xb = *reinterpret_cast<const uchar4*>(cj + s0);
        const double w0 = w[s0], w1 = w[s0 + 1], w2 = w[s0 + 2], w3 = w[s0 + 3];
        // pad samples carry state GAP and weight 0 -> add 0 to cell (gap,gap)
        bins[(xb.x * 5 + xa.x) * 32 + lane] += w0;
        bins[(xb.y * 5 + xa.y) * 32 + lane] += w1;
        bins[(xb.z * 5 + xa.z) * 32 + lane] += w2;
        bins[(xb.w * 5 + xa.w) * 32 + lane] += w3;
    }
    __syncwarp();
    double c = 0.0;
    if (lane < 25) {
#pragma unroll 8
        for (int k = 0; k < 32; ++k) c += bins[lane * 32 + ((k + lane) & 31)];
    }
    __syncwarp();
    return c;
}

// One pass of normalize_and_get_mi_single over the lane-distributed table (lane l<25: b=l/5,
// a=l%5). v is updated in place exactly like the reference's buffer (pseudocount added and the
// rows of present j-states divided by the normalisation constant), so that a second call with
// the gap bits removed reproduces the reference's MI-without-gaps quirk (mi.hpp:108-135,465-489).
__device__ __forceinline__ double warp_mi_pass(double& v, uint32_t mask_i, uint32_t mask_j, double pc, uint32_t lane) {
    const bool valid = lane < 25;
    const uint32_t a = lane % 5, b = lane / 5;
    const bool in_i = valid && ((mask_i >> a) & 1u);
    const bool in_j = valid && ((mask_j >> b) & 1u);
    if (in_i && in_j) v += pc;
    const double Z = warp_sum((in_i && in_j) ? v : 0.0);
    if (in_j) v = v / Z;
    const double jointH = warp_sum((in_i && in_j) ? v * log(v) : 0.0);
    // row sums over all five i-states (mi.hpp:173) for rows b in P_j; column sums over rows in P_j
    double rs = 0.0, cs = 0.0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        const double r = __shfl_sync(0xffffffffu, v, (lane < 5 ? lane * 5 + k : 0));
        const double q = __shfl_sync(0xffffffffu, v, (lane < 5 ? k * 5 + lane : 0));
        rs += r;
        if ((mask_j >> k) & 1u) cs += q;
    }
    const bool l5 = lane < 5;
    const double icondH = warp_sum((l5 && ((mask_j >> lane) & 1u)) ? rs * log(rs) : 0.0);
    const double jcondH = warp_sum((l5 && ((mask_i >> lane) & 1u)) ? cs * log(cs) : 0.0);
    return jointH - icondH - jcondH;
}

// =============================================================================================
// Weight-grouped popcount crosstable. Sample weights take few distinct values (w = 1/#neighbours, or all
// 1): ordering the samples by weight value and padding each group to whole words turns the weighted
// joint count into  N_ab = sum_w gval[w] * popc(Bi_a[w] & Bj_b[w])  -- ~30x fewer instructions per pair
// than scattering per-sample weights into 25 bins. Only the kept x kept cells (all present states but the
// highest-coded one) are accumulated; the remaining row / column / corner follow from the fp64 marginals
// (summed the same way), exactly as a crosstable's margins do.
// =============================================================================================
__global__ void __launch_bounds__(128)
gbits_build_kernel(const uint8_t* __restrict__ colmajor, const int32_t* __restrict__ gslot, const double* __restrict__ gval,
                   uint32_t* __restrict__ gbits, double* __restrict__ gmarg, uint32_t L, uint32_t n_pad, uint32_t NW, uint32_t col0) {
    // one CTA per column; a warp builds one 32-sample word per iteration: lane b gathers the state of slot 32w+b
    // (the column's n_pad bytes are L1-resident) and five ballots give the five plane words, staged in shared
    // memory so that the column's 5 x NW words leave in one contiguous, coalesced burst.
    extern __shared__ uint32_t planes_sm[];                  // [NSTATES][NW]
    __shared__ double red[4][NSTATES];
    const uint32_t c = col0 + blockIdx.x, lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint8_t* col = colmajor + (size_t)c * n_pad;
    for (uint32_t w = wid; w < NW; w += 4) {
        const int32_t sidx = gslot[w * 32 + lane];
        const uint32_t x = (c < L && sidx >= 0) ? (uint32_t)col[sidx] : 255u;
        // four votes; the fifth plane is what is left of the occupied slots
        const uint32_t p0 = __ballot_sync(0xffffffffu, x == 0u), p1 = __ballot_sync(0xffffffffu, x == 1u);
        const uint32_t p2 = __ballot_sync(0xffffffffu, x == 2u), p3 = __ballot_sync(0xffffffffu, x == 3u);
        const uint32_t occ = __ballot_sync(0xffffffffu, x < (uint32_t)NSTATES);
        if (lane == 0) {                                     // five predicated stores by one lane: the per-lane select of the five
            planes_sm[w] = p0; planes_sm[NW + w] = p1;        // ballots compiled into a branch ladder twice as long as the votes
            planes_sm[2 * NW + w] = p2; planes_sm[3 * NW + w] = p3;
            planes_sm[4 * NW + w] = occ & ~(p0 | p1 | p2 | p3);
        }
    }
    __syncthreads();
    // marginals from the staged planes: warp `wid` sums its share of the words, lane a of the warp ends with state a
    double m[NSTATES] = {0, 0, 0, 0, 0};
    for (uint32_t w = threadIdx.x; w < NW; w += blockDim.x) {
        const double v = gval[w];
#pragma unroll
        for (int a = 0; a < NSTATES; ++a) m[a] += v * (double)__popc(planes_sm[a * NW + w]);
    }
#pragma unroll
    for (int a = 0; a < NSTATES; ++a) {
        for (int o = 16; o; o >>= 1) m[a] += __shfl_xor_sync(0xffffffffu, m[a], o);
        if (lane == 0) red[wid][a] = m[a];
    }
    __syncthreads();
    uint32_t* out = gbits + (size_t)c * NSTATES * NW;
    for (uint32_t k = threadIdx.x; k < NSTATES * NW; k += blockDim.x) out[k] = planes_sm[k];
    if (threadIdx.x < NSTATES) gmarg[(size_t)c * NSTATES + threadIdx.x] = red[0][threadIdx.x] + red[1][threadIdx.x] + red[2][threadIdx.x] + red[3][threadIdx.x];
}

template <int KI, int KJ>
__device__ __forceinline__ void gacc(const uint32_t* __restrict__ Bi, const uint32_t* __restrict__ Bj, const uint32_t* si, const uint32_t* sj,
                                     const double* __restrict__ gval, uint32_t NW, uint32_t lane, double* acc /*[16]*/) {
    double t[KI * KJ];
#pragma unroll
    for (int k = 0; k < KI * KJ; ++k) t[k] = 0.0;
    for (uint32_t w = lane; w < NW; w += 32) {
        const double v = gval[w];
        uint32_t xi[KI], xj[KJ];
#pragma unroll
        for (int a = 0; a < KI; ++a) xi[a] = Bi[si[a] * NW + w];
#pragma unroll
        for (int b = 0; b < KJ; ++b) xj[b] = Bj[sj[b] * NW + w];
#pragma unroll
        for (int a = 0; a < KI; ++a)
#pragma unroll
            for (int b = 0; b < KJ; ++b) t[a * KJ + b] = fma((double)__popc(xi[a] & xj[b]), v, t[a * KJ + b]);
    }
#pragma unroll
    for (int a = 0; a < KI; ++a)
#pragma unroll
        for (int b = 0; b < KJ; ++b) acc[a * 4 + b] = warp_sum(t[a * KJ + b]);
}

// Returns the cell value for lane < 25 (b = lane/5 j-state, a = lane%5 i-state), like warp_crosstab.
__device__ __forceinline__ double warp_crosstab_grouped(const uint32_t* __restrict__ gbits, const double* __restrict__ gval,
                                                        const double* __restrict__ gmarg, uint32_t NW, uint32_t i, uint32_t j,
                                                        uint32_t mask_i, uint32_t mask_j, uint32_t lane) {
    // kept states = present states minus the highest-coded present one
    uint32_t si[4] = {0, 0, 0, 0}, sj[4] = {0, 0, 0, 0};
    const uint32_t di = 31u - __clz(mask_i | 1u), dj = 31u - __clz(mask_j | 1u);      // dropped (recovered) states
    uint32_t mk = mask_i & ~(1u << di), ki = 0, kj = 0;
#pragma unroll
    for (int a = 0; a < 4; ++a) if (mk) { si[a] = __ffs(mk) - 1; mk &= mk - 1; ++ki; }
    mk = mask_j & ~(1u << dj);
#pragma unroll
    for (int b = 0; b < 4; ++b) if (mk) { sj[b] = __ffs(mk) - 1; mk &= mk - 1; ++kj; }
    const uint32_t* Bi = gbits + (size_t)i * NSTATES * NW;
    const uint32_t* Bj = gbits + (size_t)j * NSTATES * NW;
    double acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = 0.0;
    // classes of 1, 2 or 4 accumulated rows; a 3-row column borrows the (all-zero) plane of an absent state
    if (ki == 3) si[3] = __ffs(~mask_i & 31u) - 1;
    if (kj == 3) sj[3] = __ffs(~mask_j & 31u) - 1;
    const uint32_t ci = ki == 3 ? 4u : ki, cj = kj == 3 ? 4u : kj;
    switch (ci * 5 + cj) {
#define SPK_GCASE(A, B) case (A) * 5 + (B): gacc<A, B>(Bi, Bj, si, sj, gval, NW, lane, acc); break;
        SPK_GCASE(1, 1) SPK_GCASE(1, 2) SPK_GCASE(1, 4)
        SPK_GCASE(2, 1) SPK_GCASE(2, 2) SPK_GCASE(2, 4)
        SPK_GCASE(4, 1) SPK_GCASE(4, 2) SPK_GCASE(4, 4)
#undef SPK_GCASE
        default: break;                                              // a monomorphic column: margins only
    }
    // assemble this lane's cell of the 5x5 table
    const uint32_t a = lane % 5, b = lane / 5;
    double v = 0.0;
    if (lane < 25 && ((mask_i >> a) & 1u) && ((mask_j >> b) & 1u)) {
        int ia = -1, jb = -1;                                        // kept index, -1 = dropped state
#pragma unroll
        for (int k = 0; k < 4; ++k) { if ((uint32_t)k < ki && si[k] == a) ia = k; if ((uint32_t)k < kj && sj[k] == b) jb = k; }
        const double* mi_ = gmarg + (size_t)i * NSTATES;
        const double* mj_ = gmarg + (size_t)j * NSTATES;
        double rowsum = 0.0, colsum = 0.0, tot = 0.0, mjk = 0.0;    // over kept x kept cells
#pragma unroll
        for (int x = 0; x < 4; ++x)
#pragma unroll
            for (int y = 0; y < 4; ++y) {
                const double cxy = acc[x * 4 + y];
                if (x == ia) rowsum += cxy;
                if (y == jb) colsum += cxy;
                tot += cxy;
            }
#pragma unroll
        for (int y = 0; y < 4; ++y) if ((uint32_t)y < kj) mjk += mj_[sj[y]];
        if (ia >= 0 && jb >= 0) {
            double cell = 0.0;
#pragma unroll
            for (int x = 0; x < 4; ++x)
#pragma unroll
                for (int y = 0; y < 4; ++y) if (x == ia && y == jb) cell = acc[x * 4 + y];
            v = cell;
        } else if (ia >= 0) v = mi_[a] - rowsum;                     // (kept a, dropped j-state)
        else if (jb >= 0) v = mj_[b] - colsum;                       // (dropped i-state, kept b)
        else v = mi_[di] - (mjk - tot);                              // corner
        v = fmax(v, 0.0);
    }
    return v;
}

// marginal-entropy tables for the grouped path: gH[c][m-1] = sum_{a in P_c} p_a ln p_a with
// p_a = (marg_a + pc*m) / (W + pc*|P_c|*m), m = #present states of the partner column. These are the row /
// column sums of the normalised joint table (mi.hpp:173-178) in closed form.
__global__ void gh_tables_kernel(const double* __restrict__ gmarg, const uint8_t* __restrict__ masks, double* __restrict__ gH,
                                 uint32_t L_pad, double w_total, double pc) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= L_pad) return;
    const uint32_t mask = masks[c] & 31u;
    const int np = __popc(mask);
    for (int m = 1; m <= NSTATES; ++m) {
        const double Z = w_total + pc * (double)np * (double)m;
        double h = 0.0;
        for (int a = 0; a < NSTATES; ++a) if ((mask >> a) & 1u) { const double pr = (gmarg[(size_t)c * NSTATES + a] + pc * (double)m) / Z; h += pr * log(pr); }
        gH[(size_t)c * NSTATES + (m - 1)] = h;
    }
}

// MI of one pair on the grouped path. Pass 1 needs one log per lane and one reduction (marginal terms come from
// gH); the normalised table is left in v exactly as the reference leaves its buffer, so that the literal
// "MI without gaps" second pass (mi.hpp:108-135,465-489) can run on it when the caller needs it.
__device__ __forceinline__ double grouped_mi_pass1(double& v, uint32_t mask_i, uint32_t mask_j, double w_total, double pc,
                                                   const double* __restrict__ gH, uint32_t i, uint32_t j, uint32_t lane) {
    const uint32_t a = lane % 5, b = lane / 5;
    const bool in_ij = lane < 25 && ((mask_i >> a) & 1u) && ((mask_j >> b) & 1u);
    const int np_i = __popc(mask_i), np_j = __popc(mask_j);
    const double Z = w_total + pc * (double)(np_i * np_j);
    double t = 0.0;
    if (in_ij) { v = (v + pc) / Z; t = v * log(v); } else v = 0.0;
    const double jointH = warp_sum(t);
    return jointH - gH[(size_t)i * NSTATES + (np_j - 1)] - gH[(size_t)j * NSTATES + (np_i - 1)];
}

__global__ void __launch_bounds__(EX_WARPS * 32, 4)
mi_pairs_grouped_kernel(const uint8_t* __restrict__ masks, const uint32_t* __restrict__ pi, const uint32_t* __restrict__ pj,
                        unsigned long long npairs, double pc, double w_total, double* __restrict__ mi,
                        const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg,
                        const double* __restrict__ gH, uint32_t NW) {
    const uint32_t lane = threadIdx.x % 32, wid = threadIdx.x / 32;
    for (unsigned long long p = (unsigned long long)blockIdx.x * EX_WARPS + wid; p < npairs;
         p += (unsigned long long)gridDim.x * EX_WARPS) {
        const uint32_t i = pi[p], j = pj[p];
        const uint32_t mi_ = masks[i], mj_ = masks[j];
        double v = warp_crosstab_grouped(gbits, gval, gmarg, NW, i, j, mi_, mj_, lane);
        const double m = grouped_mi_pass1(v, mi_, mj_, w_total, pc, gH, i, j, lane);
        if (lane == 0) mi[p] = m;
    }
}

__global__ void __launch_bounds__(EX_WARPS * 32, 4)
mi_refine_grouped_kernel(const uint8_t* __restrict__ masks, const Candidate* __restrict__ cand, unsigned long long ncand,
                         double pc, double w_total, double wo_threshold,
                         spk_edge* __restrict__ refined, double* __restrict__ refined_wo, unsigned long long* __restrict__ counters,
                         const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg,
                         const double* __restrict__ gH, uint32_t NW) {
    const uint32_t lane = threadIdx.x % 32, wid = threadIdx.x / 32;
    double maxerr = 0.0;
    for (unsigned long long p = (unsigned long long)blockIdx.x * EX_WARPS + wid; p < ncand;
         p += (unsigned long long)gridDim.x * EX_WARPS) {
        const Candidate cd = cand[p];
        const uint32_t mi_ = masks[cd.i], mj_ = masks[cd.j];
        double v = warp_crosstab_grouped(gbits, gval, gmarg, NW, cd.i, cd.j, mi_, mj_, lane);
        const double m = grouped_mi_pass1(v, mi_, mj_, w_total, pc, gH, cd.i, cd.j, lane);
        double m2 = __longlong_as_double(0x7ff8000000000000ll);   // NaN = "no gap column involved" / not needed
        // the gap-free value is only reported for SAVED edges (mi.hpp:465): skip the second pass otherwise
        if (((mi_ | mj_) & (1u << GAP)) && m > wo_threshold) m2 = warp_mi_pass(v, mi_ & ~(1u << GAP), mj_ & ~(1u << GAP), pc, lane);
        if (lane == 0) {
            spk_edge e; e.i = cd.i; e.j = cd.j; e.mi = m;
            refined[p] = e; refined_wo[p] = m2;
            if (cd.mi32 == cd.mi32) maxerr = fmax(maxerr, fabs((double)cd.mi32 - m));
        }
    }
    if (lane == 0 && maxerr > 0.0) atomicMax(&counters[3], f64_ordered(maxerr));
}

// =============================================================================================
// Thread-per-pair evaluation on the grouped bit planes. A warp-per-pair layout leaves most lanes idle in the
// entropy step (a biallelic pair has 9 cells) and spends its time in shuffle reductions; here every thread owns
// one pair: popcount-accumulates its KI x KJ kept cells over the plane words (16-byte loads), completes the
// table from the fp64 margins and evaluates the reference's formula sequentially. Candidates arrive grouped by
// K2 tile, so the threads of a warp mostly share the class pair (little divergence) and one of the two columns
// (broadcast loads).
// =============================================================================================
__device__ __forceinline__ double popc_f64(uint32_t x) {
    return __hiloint2double(0x43300000, __popc(x)) - 4503599627370496.0;
}

// kept one-hot rows of a column pair: si / sj = plane (state) of each kept row (an absent state's all-zero plane
// beyond ki / kj), di / dj = the recovered state (highest present: the gap if the column has gaps)
template <int KI, int KJ>
struct PairRows { uint32_t si[KI], sj[KJ], ki, kj, di, dj; };

template <int KI, int KJ>
__device__ __forceinline__ PairRows<KI, KJ> pair_rows(uint32_t mask_i, uint32_t mask_j) {
    PairRows<KI, KJ> R;
    uint32_t* si = R.si; uint32_t* sj = R.sj;
    const uint32_t di = 31u - __clz(mask_i | 1u), dj = 31u - __clz(mask_j | 1u);      // recovered states (gap if present)
    const uint32_t zi = __ffs(~mask_i & 31u) - 1, zj = __ffs(~mask_j & 31u) - 1;      // an absent state: all-zero plane
    uint32_t mk = mask_i & ~(1u << di), ki = 0, kj = 0;
#pragma unroll
    for (int a = 0; a < KI; ++a) { if (mk) { si[a] = __ffs(mk) - 1; mk &= mk - 1; ++ki; } else si[a] = zi; }
    mk = mask_j & ~(1u << dj);
#pragma unroll
    for (int b = 0; b < KJ; ++b) { if (mk) { sj[b] = __ffs(mk) - 1; mk &= mk - 1; ++kj; } else sj[b] = zj; }
    R.ki = ki; R.kj = kj; R.di = di; R.dj = dj;
    return R;
}

// from the KI x KJ kept cells to MI (and, for saved pairs touching a gap column, the literal MI-without-gaps)
template <int KI, int KJ>
__device__ __forceinline__ void pair_tail(const double (&acc)[KI][KJ], const PairRows<KI, KJ>& R,
                                       const double* __restrict__ gmarg, const double* __restrict__ gH,
                                       uint32_t i, uint32_t j, uint32_t mask_i, uint32_t mask_j, double w_total, double pc,
                                       double wo_threshold, double* out_m, double* out_m2) {
    const uint32_t* si = R.si; const uint32_t* sj = R.sj;
    const uint32_t ki = R.ki, kj = R.kj, di = R.di;
    // full table over (kept rows + recovered row) x (kept cols + recovered col); index KI / KJ = recovered state
    double N[KI + 1][KJ + 1];
    double colsum[KJ];
#pragma unroll
    for (int b = 0; b < KJ; ++b) colsum[b] = 0.0;
    double dsum = 0.0;
#pragma unroll
    for (int a = 0; a < KI; ++a) {
        double rs = 0.0;
#pragma unroll
        for (int b = 0; b < KJ; ++b) { N[a][b] = acc[a][b]; rs += acc[a][b]; colsum[b] += acc[a][b]; }
        const double ma = ((uint32_t)a < ki) ? gmarg[(size_t)i * NSTATES + si[a]] : 0.0;
        N[a][KJ] = fmax(ma - rs, 0.0);
    }
#pragma unroll
    for (int b = 0; b < KJ; ++b) {
        const double mb = ((uint32_t)b < kj) ? gmarg[(size_t)j * NSTATES + sj[b]] : 0.0;
        N[KI][b] = fmax(mb - colsum[b], 0.0);
        dsum += N[KI][b];
    }
    N[KI][KJ] = fmax(gmarg[(size_t)i * NSTATES + di] - dsum, 0.0);
    // ---- pass 1 (mi.hpp:146-181): p = (N + pc)/Z over present x present cells
    const int np_i = (int)ki + 1, np_j = (int)kj + 1;
    const double Z = w_total + pc * (double)(np_i * np_j);
    double jointH = 0.0;
#pragma unroll
    for (int a = 0; a <= KI; ++a)
#pragma unroll
        for (int b = 0; b <= KJ; ++b) {
            const bool valid = (a == KI || (uint32_t)a < ki) && (b == KJ || (uint32_t)b < kj);
            const double pr = (N[a][b] + pc) / Z;
            N[a][b] = valid ? pr : 0.0;
            jointH += valid ? pr * log(pr) : 0.0;
        }
    const double m = jointH - gH[(size_t)i * NSTATES + (np_j - 1)] - gH[(size_t)j * NSTATES + (np_i - 1)];
    *out_m = m;
    double m2 = __longlong_as_double(0x7ff8000000000000ll);          // NaN = no gap column involved / not needed
    const bool gi = (mask_i >> GAP) & 1u, gj = (mask_j >> GAP) & 1u;
    if ((gi || gj) && m > wo_threshold) {
        // ---- pass 2, literal (mi.hpp:108-135,465-489): the already normalised table, gap states removed from
        // the masks, pseudocount added and normalised again; row sums still run over all five i-states.
        double Z2 = 0.0;
#pragma unroll
        for (int a = 0; a <= KI; ++a)
#pragma unroll
            for (int b = 0; b <= KJ; ++b) {
                const bool in2 = ((a == KI) ? !gi : ((uint32_t)a < ki)) && ((b == KJ) ? !gj : ((uint32_t)b < kj));
                if (in2) { N[a][b] += pc; Z2 += N[a][b]; }
            }
        double jH = 0.0, cH = 0.0;
        double rs2[KJ + 1];
#pragma unroll
        for (int b = 0; b <= KJ; ++b) rs2[b] = 0.0;
#pragma unroll
        for (int a = 0; a <= KI; ++a) {
            const bool ain = (a == KI) ? !gi : ((uint32_t)a < ki);
            double cs = 0.0;
#pragma unroll
            for (int b = 0; b <= KJ; ++b) {
                const bool bin = (b == KJ) ? !gj : ((uint32_t)b < kj);
                if (bin) {
                    const double q = N[a][b] / Z2;               // every i-state of a present j row is divided (mi.hpp:165)
                    rs2[b] += q;                                  // row sum over ALL i-states, the gap_i cell included
                    if (ain) { jH += q * log(q); cs += q; }
                }
            }
            if (ain) cH += cs * log(cs);
        }
        double rH = 0.0;
#pragma unroll
        for (int b = 0; b <= KJ; ++b) {
            const bool bin = (b == KJ) ? !gj : ((uint32_t)b < kj);
            if (bin) rH += rs2[b] * log(rs2[b]);
        }
        m2 = jH - rH - cH;
    }
    *out_m2 = m2;
}

template <int KI, int KJ>
__device__ __forceinline__ void pair_eval(const uint32_t* __restrict__ gbits, const double* __restrict__ gval,
                                       const double* __restrict__ gmarg, const double* __restrict__ gH, uint32_t NW,
                                       uint32_t i, uint32_t j, uint32_t mask_i, uint32_t mask_j, double w_total, double pc,
                                       double wo_threshold, double* out_m, double* out_m2) {
    const PairRows<KI, KJ> R = pair_rows<KI, KJ>(mask_i, mask_j);
    const uint32_t* si = R.si; const uint32_t* sj = R.sj;
    const uint4* Bi[KI]; const uint4* Bj[KJ];
#pragma unroll
    for (int a = 0; a < KI; ++a) Bi[a] = reinterpret_cast<const uint4*>(gbits + ((size_t)i * NSTATES + si[a]) * NW);
#pragma unroll
    for (int b = 0; b < KJ; ++b) Bj[b] = reinterpret_cast<const uint4*>(gbits + ((size_t)j * NSTATES + sj[b]) * NW);
    double acc[KI][KJ];
#pragma unroll
    for (int a = 0; a < KI; ++a)
#pragma unroll
        for (int b = 0; b < KJ; ++b) acc[a][b] = 0.0;
    const double2* gv2 = reinterpret_cast<const double2*>(gval);
    for (uint32_t w4 = 0; w4 < NW / 4; ++w4) {
        const double2 v01 = gv2[2 * w4], v23 = gv2[2 * w4 + 1];
        uint4 xi[KI], xj[KJ];
#pragma unroll
        for (int a = 0; a < KI; ++a) xi[a] = __ldg(Bi[a] + w4);
#pragma unroll
        for (int b = 0; b < KJ; ++b) xj[b] = __ldg(Bj[b] + w4);
#pragma unroll
        for (int a = 0; a < KI; ++a)
#pragma unroll
            for (int b = 0; b < KJ; ++b) {
                // popcount -> double without I2F.F64 (which shares the quarter-rate pipe with POPC): 2^52 + c, minus 2^52 (exact)
                double t = acc[a][b];
                t = fma(popc_f64(xi[a].x & xj[b].x), v01.x, t);
                t = fma(popc_f64(xi[a].y & xj[b].y), v01.y, t);
                t = fma(popc_f64(xi[a].z & xj[b].z), v23.x, t);
                t = fma(popc_f64(xi[a].w & xj[b].w), v23.y, t);
                acc[a][b] = t;
            }
    }
    pair_tail<KI, KJ>(acc, R, gmarg, gH, i, j, mask_i, mask_j, w_total, pc, wo_threshold, out_m, out_m2);
}

template <int KI, int KJ>
__device__ __noinline__ void pair_eval_nl(const uint32_t* __restrict__ gbits, const double* __restrict__ gval,
                                          const double* __restrict__ gmarg, const double* __restrict__ gH, uint32_t NW,
                                          uint32_t i, uint32_t j, uint32_t mask_i, uint32_t mask_j, double w_total, double pc,
                                          double wo_threshold, double* out_m, double* out_m2) {
    pair_eval<KI, KJ>(gbits, gval, gmarg, gH, NW, i, j, mask_i, mask_j, w_total, pc, wo_threshold, out_m, out_m2);
}

__device__ __forceinline__ void pair_eval_dispatch(const uint32_t* __restrict__ gbits, const double* __restrict__ gval,
                                                   const double* __restrict__ gmarg, const double* __restrict__ gH, uint32_t NW,
                                                   uint32_t i, uint32_t j, uint32_t mask_i, uint32_t mask_j, double w_total, double pc,
                                                   double wo_threshold, double* m, double* m2) {
    const int ki = __popc(mask_i) - 1, kj = __popc(mask_j) - 1;
    const int ci = ki <= 1 ? 1 : (ki == 2 ? 2 : 4), cj = kj <= 1 ? 1 : (kj == 2 ? 2 : 4);
    switch (ci * 8 + cj) {
#define SPK_PCASE(A, B) case (A) * 8 + (B): pair_eval_nl<A, B>(gbits, gval, gmarg, gH, NW, i, j, mask_i, mask_j, w_total, pc, wo_threshold, m, m2); break;
        SPK_PCASE(1, 1) SPK_PCASE(1, 2) SPK_PCASE(1, 4)
        SPK_PCASE(2, 1) SPK_PCASE(2, 2) SPK_PCASE(2, 4)
        SPK_PCASE(4, 1) SPK_PCASE(4, 2) SPK_PCASE(4, 4)
#undef SPK_PCASE
        default: break;
    }
}

__global__ void __launch_bounds__(128)
mi_pairs_thread_kernel(const uint8_t* __restrict__ masks, const uint32_t* __restrict__ pi, const uint32_t* __restrict__ pj,
                       unsigned long long npairs, double pc, double w_total, double* __restrict__ mi,
                       const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg,
                       const double* __restrict__ gH, uint32_t NW) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const uint32_t i = pi[p], j = pj[p];
    double m = 0.0, m2 = 0.0;
    pair_eval_dispatch(gbits, gval, gmarg, gH, NW, i, j, masks[i], masks[j], w_total, pc, INFINITY, &m, &m2);
    mi[p] = m;
}

__global__ void __launch_bounds__(128)
mi_refine_thread_kernel(const uint8_t* __restrict__ masks, const Candidate* __restrict__ cand, unsigned long long ncand,
                        double pc, double w_total, double wo_threshold,
                        spk_edge* __restrict__ refined, double* __restrict__ refined_wo, unsigned long long* __restrict__ counters,
                        const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg,
                        const double* __restrict__ gH, uint32_t NW) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    double err = 0.0;
    if (p < ncand) {
        const Candidate cd = cand[p];
        double m = 0.0, m2 = 0.0;
        pair_eval_dispatch(gbits, gval, gmarg, gH, NW, cd.i, cd.j, masks[cd.i], masks[cd.j], w_total, pc, wo_threshold, &m, &m2);
        spk_edge e; e.i = cd.i; e.j = cd.j; e.mi = m;
        refined[p] = e; refined_wo[p] = m2;
        if (cd.mi32 == cd.mi32) err = fabs((double)cd.mi32 - m);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) err = fmax(err, __shfl_xor_sync(0xffffffffu, err, o));
    if ((threadIdx.x & 31) == 0 && err > 0.0 && f64_ordered(err) > *(volatile unsigned long long*)&counters[3]) atomicMax(&counters[3], f64_ordered(err));
}

// ---- class-sorted refine: candidates are bucketed by the (rows_i, rows_j) class pair so that every launch runs
// ONE instantiation with converged warps (the 9-way divergent dispatch thrashes the instruction cache).
__device__ __forceinline__ uint32_t pair_class(uint32_t mask_i, uint32_t mask_j) {
    const int ki = __popc(mask_i) - 1, kj = __popc(mask_j) - 1;
    return (uint32_t)((ki <= 1 ? 0 : (ki == 2 ? 1 : 2)) * 3 + (kj <= 1 ? 0 : (kj == 2 ? 1 : 2)));
}
__global__ void __launch_bounds__(256)
classify_count_kernel(const Candidate* __restrict__ cand, unsigned long long ncand, const uint8_t* __restrict__ masks,
                      unsigned long long* __restrict__ counts /*[9]*/) {
    __shared__ unsigned int h[9];
    if (threadIdx.x < 9) h[threadIdx.x] = 0;
    __syncthreads();
    for (unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; p < ncand; p += (unsigned long long)gridDim.x * blockDim.x) {
        const Candidate cd = cand[p];
        if (cd.i == I8_INVALID) continue;                          // unused slot of an emission chunk (int8 engine)
        atomicAdd(&h[pair_class(masks[cd.i], masks[cd.j])], 1u);
    }
    __syncthreads();
    if (threadIdx.x < 9 && h[threadIdx.x]) atomicAdd(&counts[threadIdx.x], (unsigned long long)h[threadIdx.x]);
}
__global__ void __launch_bounds__(256)
classify_scatter_kernel(const Candidate* __restrict__ cand, unsigned long long ncand, const uint8_t* __restrict__ masks,
                        unsigned long long* __restrict__ cursors /*[9], preset to the class offsets*/, Candidate* __restrict__ out) {
    __shared__ unsigned int h[9];
    __shared__ unsigned long long base[9];
    const unsigned long long chunk = (unsigned long long)blockDim.x * 8;
    for (unsigned long long c0 = (unsigned long long)blockIdx.x * chunk; c0 < ncand; c0 += (unsigned long long)gridDim.x * chunk) {
        if (threadIdx.x < 9) h[threadIdx.x] = 0;
        __syncthreads();
        Candidate cd[8]; uint32_t cls[8], loc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const unsigned long long p = c0 + (unsigned long long)k * blockDim.x + threadIdx.x;
            cls[k] = 9;
            if (p < ncand) {
                cd[k] = cand[p];
                if (cd[k].i != I8_INVALID) { cls[k] = pair_class(masks[cd[k].i], masks[cd[k].j]); loc[k] = atomicAdd(&h[cls[k]], 1u); }
            }
        }
        __syncthreads();
        if (threadIdx.x < 9) base[threadIdx.x] = h[threadIdx.x] ? atomicAdd(&cursors[threadIdx.x], (unsigned long long)h[threadIdx.x]) : 0ull;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 8; ++k) if (cls[k] < 9) out[base[cls[k]] + loc[k]] = cd[k];
        __syncthreads();
    }
}

template <int KI, int KJ>
__global__ void __launch_bounds__(128)
mi_refine_class_kernel(const uint8_t* __restrict__ masks, const Candidate* __restrict__ cand, unsigned long long ncand,
                       double pc, double w_total, double wo_threshold,
                       spk_edge* __restrict__ refined, double* __restrict__ refined_wo, unsigned long long* __restrict__ counters,
                       const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg,
                       const double* __restrict__ gH, uint32_t NW) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    double err = 0.0;
    if (p < ncand) {
        const Candidate cd = cand[p];
        double m = 0.0, m2 = 0.0;
        pair_eval<KI, KJ>(gbits, gval, gmarg, gH, NW, cd.i, cd.j, masks[cd.i], masks[cd.j], w_total, pc, wo_threshold, &m, &m2);
        spk_edge e; e.i = cd.i; e.j = cd.j; e.mi = m;
        refined[p] = e; refined_wo[p] = m2;
        if (cd.mi32 == cd.mi32) err = fabs((double)cd.mi32 - m);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) err = fmax(err, __shfl_xor_sync(0xffffffffu, err, o));
    if ((threadIdx.x & 31) == 0 && err > 0.0 && f64_ordered(err) > *(volatile unsigned long long*)&counters[3]) atomicMax(&counters[3], f64_ordered(err));
}

// Cooperative version of mi_refine_class_kernel. Thread-per-pair plane loads are 32 scattered 16-byte requests per
// instruction: ncu showed the class kernels L1TEX-bound (l1tex throughput 86-96 %, issue slots 18-36 % busy). Here a
// group of G lanes walks the planes of ONE pair together (G x 16 contiguous bytes per load: whole 128-byte lines for
// G = 8), one pair of the group after the other; the G partial tables of the group's G pairs are then summed with a
// transposing butterfly (G-1 exchanges per cell instead of G log G) that leaves lane t of the group with the complete
// table of ITS pair, and the entropy tail runs thread-per-pair as before. G = min(8, 32 / (KI*KJ)) bounds the
// partial tables at 32 doubles per thread.
#ifndef SPK_REFINE_UNROLL
#define SPK_REFINE_UNROLL 1           // plane-walk iterations in flight per thread (A/B: tools/build_variant.sh)
#endif
constexpr int REFINE_UNROLL = SPK_REFINE_UNROLL;
#ifndef SPK_REFINE_MINBLOCKS
#define SPK_REFINE_MINBLOCKS 6           // 6 x 128 threads resident: 7.8 -> 7.0 ms on config 2 (unroll 2 / 5 were slower)
#endif
template <int KI, int KJ, int G>
__global__ void __launch_bounds__(128, SPK_REFINE_MINBLOCKS)
mi_refine_coop_kernel(const uint8_t* __restrict__ masks, const Candidate* __restrict__ cand, unsigned long long ncand,
                      double pc, double w_total, double wo_threshold,
                      spk_edge* __restrict__ refined, double* __restrict__ refined_wo, unsigned long long* __restrict__ counters,
                      const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg,
                      const double* __restrict__ gH, uint32_t NW) {
    constexpr int C = KI * KJ;
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x % 32, sub = lane % G;
    const bool live = p < ncand;
    Candidate cd; cd.i = 0; cd.j = 0; cd.mi32 = 0.f;
    if (live) cd = cand[p];
    const uint32_t mask_i = masks[cd.i], mask_j = masks[cd.j];
    const PairRows<KI, KJ> R = pair_rows<KI, KJ>(mask_i, mask_j);
    uint32_t rowi[KI], rowj[KJ];                    // plane row index (column * NSTATES + state)
#pragma unroll
    for (int a = 0; a < KI; ++a) rowi[a] = cd.i * NSTATES + R.si[a];
#pragma unroll
    for (int b = 0; b < KJ; ++b) rowj[b] = cd.j * NSTATES + R.sj[b];
    const double2* gv2 = reinterpret_cast<const double2*>(gval);
    const uint32_t nchunk = NW / 4;                  // uint4 chunks per plane row (NW is a multiple of 32)

    // Partial tables are folded as soon as a sibling is complete (a binary counter: after pair t, one fold per trailing one
    // bit of t), so at most log2(G) folded tables are alive instead of G partial ones. The fold at distance h keeps the
    // lower sibling in lanes whose bit h is clear and the upper one in the others: lane t of the group ends with pair t.
    constexpr int LOG2G = (G == 8) ? 3 : (G == 4 ? 2 : 1);
    double stack[LOG2G][C];
    double acc2[KI][KJ];
#pragma unroll
    for (int t = 0; t < G; ++t) {
        const uint4* Bi[KI]; const uint4* Bj[KJ];
#pragma unroll
        for (int a = 0; a < KI; ++a) Bi[a] = reinterpret_cast<const uint4*>(gbits + (size_t)__shfl_sync(0xffffffffu, rowi[a], t, G) * NW);
#pragma unroll
        for (int b = 0; b < KJ; ++b) Bj[b] = reinterpret_cast<const uint4*>(gbits + (size_t)__shfl_sync(0xffffffffu, rowj[b], t, G) * NW);
        double acc[C];
#pragma unroll
        for (int q = 0; q < C; ++q) acc[q] = 0.0;
#pragma unroll (REFINE_UNROLL)
        for (uint32_t w4 = sub; w4 < nchunk; w4 += G) {
            const double2 v01 = gv2[2 * w4], v23 = gv2[2 * w4 + 1];
            uint4 xi[KI], xj[KJ];
#pragma unroll
            for (int a = 0; a < KI; ++a) xi[a] = __ldg(Bi[a] + w4);
#pragma unroll
            for (int b = 0; b < KJ; ++b) xj[b] = __ldg(Bj[b] + w4);
#pragma unroll
            for (int a = 0; a < KI; ++a)
#pragma unroll
                for (int b = 0; b < KJ; ++b) {
                    double v = acc[a * KJ + b];
                    v = fma(popc_f64(xi[a].x & xj[b].x), v01.x, v);
                    v = fma(popc_f64(xi[a].y & xj[b].y), v01.y, v);
                    v = fma(popc_f64(xi[a].z & xj[b].z), v23.x, v);
                    v = fma(popc_f64(xi[a].w & xj[b].w), v23.y, v);
                    acc[a * KJ + b] = v;
                }
        }
        int ones = 0;                                   // trailing one bits of t (compile-time after unrolling)
        while ((t >> ones) & 1) ++ones;
#pragma unroll
        for (int lvl = 0; lvl < LOG2G; ++lvl) {
            if (lvl < ones) {
                const int h = 1 << lvl;
                const bool up = (sub & h) != 0;
#pragma unroll
                for (int q = 0; q < C; ++q) {
                    const double send = up ? stack[lvl][q] : acc[q];
                    const double keep = up ? acc[q] : stack[lvl][q];
                    acc[q] = keep + __shfl_xor_sync(0xffffffffu, send, h);
                }
            }
        }
        if (ones < LOG2G) {
#pragma unroll
            for (int q = 0; q < C; ++q) stack[ones][q] = acc[q];
        } else {
#pragma unroll
            for (int a = 0; a < KI; ++a)
#pragma unroll
                for (int b = 0; b < KJ; ++b) acc2[a][b] = acc[a * KJ + b];
        }
    }
    double err = 0.0;
    if (live) {
        double m = 0.0, m2 = 0.0;
        pair_tail<KI, KJ>(acc2, R, gmarg, gH, cd.i, cd.j, mask_i, mask_j, w_total, pc, wo_threshold, &m, &m2);
        spk_edge e; e.i = cd.i; e.j = cd.j; e.mi = m;
        refined[p] = e; refined_wo[p] = m2;
        if (cd.mi32 == cd.mi32) err = fabs((double)cd.mi32 - m);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) err = fmax(err, __shfl_xor_sync(0xffffffffu, err, o));
    if (lane == 0 && err > 0.0 && f64_ordered(err) > *(volatile unsigned long long*)&counters[3]) atomicMax(&counters[3], f64_ordered(err));   // (read first: the maximum settles early, 4 x 10^5 same-address atomics otherwise)
}

// ---- K2b: explicit pair list ---------------------------------------------------------------
__global__ void __launch_bounds__(EX_WARPS * 32)
mi_pairs_kernel(const uint8_t* __restrict__ colmajor, const double* __restrict__ w, const uint8_t* __restrict__ masks,
                const uint32_t* __restrict__ pi, const uint32_t* __restrict__ pj, unsigned long long npairs,
                uint32_t n_pad, double pc, double* __restrict__ mi,
                const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg, uint32_t NW) {
    __shared__ double bins_all[EX_WARPS][25 * 32];
    const uint32_t lane = threadIdx.x % 32, wid = threadIdx.x / 32;
    double* bins = bins_all[wid];
    for (unsigned long long p = (unsigned long long)blockIdx.x * EX_WARPS + wid; p < npairs;
         p += (unsigned long long)gridDim.x * EX_WARPS) {
        const uint32_t i = pi[p], j = pj[p];
        double v = gbits ? warp_crosstab_grouped(gbits, gval, gmarg, NW, i, j, masks[i], masks[j], lane)
                         : warp_crosstab(colmajor + (size_t)i * n_pad, colmajor + (size_t)j * n_pad, w, n_pad, bins, lane);
        const double m = warp_mi_pass(v, masks[i], masks[j], pc, lane);
        if (lane == 0) mi[p] = m;
    }
}

// ---- K2r/K2c: refine filter candidates -------------------------------------------------------
// counters[3] = max |mi32 - mi64| (ordered bits, atomicMax)
__global__ void __launch_bounds__(EX_WARPS * 32)
mi_refine_kernel(const uint8_t* __restrict__ colmajor, const double* __restrict__ w, const uint8_t* __restrict__ masks,
                 const Candidate* __restrict__ cand, unsigned long long ncand, uint32_t n_pad, double pc,
                 spk_edge* __restrict__ refined, double* __restrict__ refined_wo, unsigned long long* __restrict__ counters,
                 const uint32_t* __restrict__ gbits, const double* __restrict__ gval, const double* __restrict__ gmarg, uint32_t NW) {
    __shared__ double bins_all[EX_WARPS][25 * 32];
    const uint32_t lane = threadIdx.x % 32, wid = threadIdx.x / 32;
    double* bins = bins_all[wid];
    double maxerr = 0.0;
    for (unsigned long long p = (unsigned long long)blockIdx.x * EX_WARPS + wid; p < ncand;
         p += (unsigned long long)gridDim.x * EX_WARPS) {
        const Candidate cd = cand[p];
        if (cd.i == I8_INVALID) {                                  // unused slot of an emission chunk (int8 engine): never kept
            if (lane == 0) {
                spk_edge e; e.i = 0; e.j = 0; e.mi = __longlong_as_double(0x7ff8000000000000ll);
                refined[p] = e; refined_wo[p] = e.mi;
            }
            continue;
        }
        const uint32_t mi_ = masks[cd.i], mj_ = masks[cd.j];
        double v = gbits ? warp_crosstab_grouped(gbits, gval, gmarg, NW, cd.i, cd.j, mi_, mj_, lane)
                         : warp_crosstab(colmajor + (size_t)cd.i * n_pad, colmajor + (size_t)cd.j * n_pad, w, n_pad, bins, lane);
        const double m = warp_mi_pass(v, mi_, mj_, pc, lane);
        double m2 = __longlong_as_double(0x7ff8000000000000ll);   // NaN = "no gap column involved"
        if ((mi_ | mj_) & (1u << GAP)) m2 = warp_mi_pass(v, mi_ & ~(1u << GAP), mj_ & ~(1u << GAP), pc, lane);
        if (lane == 0) {
            spk_edge e; e.i = cd.i; e.j = cd.j; e.mi = m;
            refined[p] = e; refined_wo[p] = m2;
            if (cd.mi32 == cd.mi32) maxerr = fmax(maxerr, fabs((double)cd.mi32 - m));
        }
    }
    if (lane == 0 && maxerr > 0.0) atomicMax(&counters[3], f64_ordered(maxerr));
}

// ---- filter audit ------------------------------------------------------------------------------
// The seed launch of the filter leaves an unbiased sample of pairs with mi32 > thr - 4*band in the audit buffer. A strided
// subsample of it (>= 1e5 pairs when the sample is that large) is APPENDED to the candidate list, marked pad = 1, so that the
// regular refine launches recompute it in fp64 (no kernel of its own: one thread per pair over 1e5 pairs had cost 0.84 ms, the
// cooperative refine does them in ~0.06); mi_compact_kernel turns the marked entries into the audit statistics
// (counters[8] max |mi32 - mi64|, [9] pairs with mi64 > thr that the filter would have dropped, [10] pairs audited)
// instead of edges. n is read on the device (no host round trip).
constexpr unsigned long long AUDIT_TARGET = 300000;     // buffer entries visited per scan and rank (about half of them are unused slots of the
                                                        // emission chunks: >= 1e5 pairs recomputed when the sample is that large)
__global__ void __launch_bounds__(256)
audit_append_kernel(const Candidate* __restrict__ audit, const unsigned long long* __restrict__ d_n, unsigned long long cap,
                    Candidate* __restrict__ cand, unsigned long long cand_cap, unsigned long long* __restrict__ counters) {
    const unsigned long long n = min(*d_n, cap);
    const unsigned long long stride = (n + AUDIT_TARGET - 1) / AUDIT_TARGET;     // every stride-th entry of the emission order
    for (unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; q * stride < n; q += (unsigned long long)gridDim.x * blockDim.x) {
        Candidate cd = audit[q * stride];
        if (cd.i == I8_INVALID) continue;
        const unsigned long long slot = atomicAdd(&counters[0], 1ull);            // the main pass is complete: single slots after its chunks
        if (slot < cand_cap) { cd.pad = 1u; cand[slot] = cd; } else counters[4] = 1ull;
    }
}

__global__ void __launch_bounds__(256) mi_compact_kernel(const spk_edge* __restrict__ refined, const double* __restrict__ refined_wo,
                                  unsigned long long ncand, const uint64_t* __restrict__ pos,
                                  double threshold, unsigned long long ld_threshold, int linear, unsigned long long circle,
                                  unsigned long long* __restrict__ colmax_ord,
                                  spk_edge* __restrict__ edges, unsigned long long edges_cap,
                                  spk_edge* __restrict__ wo, unsigned long long wo_cap,
                                  unsigned long long* __restrict__ counters,
                                  const Candidate* __restrict__ meta, float cand_lo) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x % 32;
    bool keep = false, keep_wo = false;
    spk_edge e; e.i = 0; e.j = 0; e.mi = 0.0; double m2 = 0.0;
    // audit entries (meta[p].pad == 1; `meta` is the candidate array `refined` runs parallel to): statistics, never edges
    double aerr = 0.0; unsigned int abad = 0, acnt = 0;
    const bool is_audit = meta && p < ncand && meta[p].pad == 1u;
    if (is_audit) {
        const double m = refined[p].mi; const float mi32 = meta[p].mi32;
        aerr = fabs((double)mi32 - m); abad = (m > threshold && !(mi32 > cand_lo)) ? 1u : 0u; acnt = 1u;
    }
    if (meta && __any_sync(0xffffffffu, is_audit)) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            aerr = fmax(aerr, __shfl_xor_sync(0xffffffffu, aerr, o));
            abad += __shfl_xor_sync(0xffffffffu, abad, o); acnt += __shfl_xor_sync(0xffffffffu, acnt, o);
        }
        if (lane == 0) {
            if (aerr > 0.0) atomicMax(&counters[8], f64_ordered(aerr));
            if (abad) atomicAdd(&counters[9], (unsigned long long)abad);
            atomicAdd(&counters[10], (unsigned long long)acnt);
        }
    }
    if (p < ncand && !is_audit) {
        e = refined[p]; m2 = refined_wo[p];
        if (genome_distance(pos[e.i], pos[e.j], linear, circle) > ld_threshold) {
            const unsigned long long o = f64_ordered(e.mi);
            if (o > colmax_ord[e.i]) atomicMax(&colmax_ord[e.i], o);
            if (o > colmax_ord[e.j]) atomicMax(&colmax_ord[e.j], o);
        }
        keep = threshold < e.mi;                 // strict (mi.hpp:430,454); NaN never kept
        keep_wo = keep && (m2 == m2);
    }
    // slots: ONE atomicAdd per block and list (per-warp reservations were 8 x 10^5 atomics on two addresses, which serialise in
    // the L2 slice that owns them: most of the kernel's 0.76 ms). blockDim.x == 256.
    __shared__ unsigned int s_cnt[2][8];
    __shared__ unsigned long long s_base[2];
    const uint32_t wid = threadIdx.x / 32;
    const uint32_t bal = __ballot_sync(0xffffffffu, keep), bal2 = __ballot_sync(0xffffffffu, keep_wo);
    if (lane == 0) { s_cnt[0][wid] = __popc(bal); s_cnt[1][wid] = __popc(bal2); }
    __syncthreads();
    if (threadIdx.x < 2) {
        unsigned int tot = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { const unsigned int v = s_cnt[threadIdx.x][w]; s_cnt[threadIdx.x][w] = tot; tot += v; }    // exclusive prefix
        s_base[threadIdx.x] = tot ? atomicAdd(&counters[1 + threadIdx.x], (unsigned long long)tot) : 0ull;
    }
    __syncthreads();
    if (keep) {
        const unsigned long long idx = s_base[0] + s_cnt[0][wid] + __popc(bal & ((1u << lane) - 1u));
        if (idx < edges_cap) edges[idx] = e; else counters[4] = 1ull;
    }
    if (keep_wo) {
        const unsigned long long idx = s_base[1] + s_cnt[1][wid] + __popc(bal2 & ((1u << lane) - 1u));
        spk_edge e2 = e; e2.mi = m2;
        if (idx < wo_cap) wo[idx] = e2; else counters[4] = 1ull;
    }
}

// ---- exhaustive engine helper: enumerate all pairs (i, j>i) for rows [i0, i1) -----------------
__global__ void enumerate_pairs_kernel(Candidate* __restrict__ cand, uint32_t i0, uint32_t i1, uint32_t L,
                                       unsigned long long total) {
    // pairs laid out row by row; row i has L-1-i entries. Closed-form inverse is avoided: one block per row.
    for (uint32_t i = i0 + blockIdx.x; i < i1; i += gridDim.x) {
        // offset of row i within [i0,i1): sum_{r=i0}^{i-1} (L-1-r)
        const unsigned long long k = (unsigned long long)(i - i0);
        const unsigned long long off = k * (unsigned long long)(L - 1 - i0) - k * (k - 1) / 2;
        for (uint32_t j = i + 1 + threadIdx.x; j < L; j += blockDim.x) {
            Candidate c; c.i = i; c.j = j; c.mi32 = __int_as_float(0x7fc00000); c.pad = 0;
            const unsigned long long idx = off + (j - i - 1);
            if (idx < total) cand[idx] = c;
        }
    }
}

// =============================================================================================
// host: group the samples by distinct weight value and build the bit planes (once per weights/alignment)
void mi_exact_free(Ctx* c) {
    if (c->d_gbits) { cudaFree(c->d_gbits); c->d_gbits = nullptr; }
    if (c->d_gval) { cudaFree(c->d_gval); c->d_gval = nullptr; }
    if (c->d_gmarg) { cudaFree(c->d_gmarg); c->d_gmarg = nullptr; }
    if (c->d_gslot) { cudaFree(c->d_gslot); c->d_gslot = nullptr; }
    if (c->d_gH) { cudaFree(c->d_gH); c->d_gH = nullptr; }
    c->g_cap_words = 0; c->g_NW = 0; c->g_state = 0; c->g_pc = -1.0; c->g_nw_cap = 0;
}

static int ensure_gh(Ctx* c, double pc) {
    if (c->g_state != 1 || c->g_pc == pc) return SPK_OK;
    if (!c->d_gH && cudaMalloc(&c->d_gH, (size_t)c->L_pad * NSTATES * 8) != cudaSuccess) {
        cudaGetLastError(); return fail(c, SPK_E_NOMEM, "grouped bit planes: allocation failed");
    }
    gh_tables_kernel<<<(c->L_pad + 127) / 128, 128, 0, c->stream>>>(c->d_gmarg, c->d_masks, c->d_gH, c->L_pad, c->w_total, pc);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    c->g_pc = pc;
    return SPK_OK;
}

static int ensure_grouped(Ctx* c, double pc) {
    if (c->g_state != 0) return ensure_gh(c, pc);
    const uint32_t n = c->n;
    if (c->h_w.size() != n) { c->g_state = 2; return SPK_OK; }
    // order samples by weight value (bit pattern), stable
    std::vector<uint32_t> order(n);
    for (uint32_t s = 0; s < n; ++s) order[s] = s;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return c->h_w[a] < c->h_w[b]; });
    // pageable sources of the async uploads below: the copies stage them before returning, but the vectors live in the context anyway
    std::vector<int32_t>& slot = c->h_gslot; std::vector<double>& val = c->h_gval;
    slot.clear(); val.clear();
    slot.reserve((size_t)n + 4096);
    for (uint32_t k = 0; k < n;) {
        uint32_t e = k;
        while (e < n && c->h_w[order[e]] == c->h_w[order[k]]) ++e;
        for (uint32_t t = k; t < e; ++t) slot.push_back((int32_t)order[t]);
        while (slot.size() % 32) slot.push_back(-1);
        while (val.size() < slot.size() / 32) val.push_back(c->h_w[order[k]]);
        k = e;
        if (slot.size() / 32 > 2 * (size_t)((n + 31) / 32) + 64) { c->g_state = 2; return SPK_OK; }   // too many distinct weights
    }
    while ((slot.size() / 32) % 32) { for (int t = 0; t < 32; ++t) slot.push_back(-1); val.push_back(0.0); }
    const uint32_t NW = (uint32_t)(slot.size() / 32);
    const size_t words = (size_t)c->L_pad * NSTATES * NW;
    if (words > c->g_cap_words || !c->d_gbits) {
        mi_exact_free(c);
        if (cudaMalloc(&c->d_gbits, words * 4) != cudaSuccess || cudaMalloc(&c->d_gmarg, (size_t)c->L_pad * NSTATES * 8) != cudaSuccess) {
            cudaGetLastError(); mi_exact_free(c);
            c->g_state = 2;                                          // not enough memory: the byte kernel still works
            return SPK_OK;
        }
        c->g_cap_words = words;
    }
    if (NW > c->g_nw_cap || !c->d_gval || !c->d_gslot) {               // grow-only (cudaFree synchronises the device)
        if (c->d_gval) { cudaFree(c->d_gval); c->d_gval = nullptr; }
        if (c->d_gslot) { cudaFree(c->d_gslot); c->d_gslot = nullptr; }
        c->g_nw_cap = 0;
        const uint32_t cap = NW + NW / 4 + 16;
        if (cudaMalloc(&c->d_gval, (size_t)cap * 8) != cudaSuccess || cudaMalloc(&c->d_gslot, (size_t)cap * 32 * 4) != cudaSuccess) {
            cudaGetLastError(); return fail(c, SPK_E_NOMEM, "grouped bit planes: allocation failed");
        }
        c->g_nw_cap = cap;
    }
    SPK_CUDA(c, cudaMemcpyAsync(c->d_gval, val.data(), (size_t)NW * 8, cudaMemcpyHostToDevice, c->stream));
    SPK_CUDA(c, cudaMemcpyAsync(c->d_gslot, slot.data(), (size_t)NW * 32 * 4, cudaMemcpyHostToDevice, c->stream));
    const size_t gsm = (size_t)NSTATES * NW * 4;
    if (gsm > 200 * 1024) { c->g_state = 2; return SPK_OK; }        // one column's planes must fit shared memory
    if (gsm > 48 * 1024) SPK_CUDA(c, cudaFuncSetAttribute(gbits_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsm));
    if (c->world > 1 && c->L_pad % (uint32_t)c->world == 0) {
        // multi-GPU: every rank builds the planes of its L_pad / G columns, the slices are all-gathered over NVLink
        // (the build depends on the weights only, so it is identical work on every rank otherwise)
        const uint32_t per = c->L_pad / (uint32_t)c->world, col0 = per * (uint32_t)c->rank;
        gbits_build_kernel<<<per, 128, gsm, c->stream>>>(c->d_colmajor, c->d_gslot, c->d_gval, c->d_gbits, c->d_gmarg, c->L, c->n_pad, NW, col0);
        int rx = comm_allgather_inplace(c, c->d_gbits, (size_t)per * NSTATES * NW * 4); if (rx) return rx;
        rx = comm_allgather_inplace(c, c->d_gmarg, (size_t)per * NSTATES * 8); if (rx) return rx;
    } else {
        gbits_build_kernel<<<c->L_pad, 128, gsm, c->stream>>>(c->d_colmajor, c->d_gslot, c->d_gval, c->d_gbits, c->d_gmarg, c->L, c->n_pad, NW, 0u);
    }
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    c->g_NW = NW; c->g_state = 1; c->g_pc = -1.0;
    return ensure_gh(c, pc);
}

static int refine_grid(Ctx* c, unsigned long long n) {
    unsigned long long blocks = (n + EX_WARPS - 1) / EX_WARPS;
    const unsigned long long cap = (unsigned long long)c->sm_count * 16;
    return (int)(blocks < cap ? (blocks ? blocks : 1) : cap);
}

int mi_refine_prepare(Ctx* c, double pc) { return ensure_grouped(c, pc); }

int mi_audit_append(Ctx* c, const Candidate* d_audit, const unsigned long long* d_n, uint64_t cap) {
    if (!d_audit || cap == 0) return SPK_OK;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AUD0], c->stream));
    audit_append_kernel<<<c->sm_count * 4, 256, 0, c->stream>>>(d_audit, d_n, cap, c->d_cand, c->cand_cap, c->d_counters);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AUD1], c->stream));
    c->tm_pending |= TMP_AUDIT;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

int mi_pairs_exact(Ctx* c, const uint32_t* d_pi, const uint32_t* d_pj, uint64_t npairs, double pc, double* d_mi) {
    if (npairs == 0) return SPK_OK;
    { int rg = ensure_grouped(c, pc); if (rg) return rg; }
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_PAIRS0], c->stream));
    if (c->g_state == 1)
        mi_pairs_thread_kernel<<<(unsigned)((npairs + 127) / 128), 128, 0, c->stream>>>(c->d_masks, d_pi, d_pj, npairs, pc, c->w_total, d_mi,
                                                                                       c->d_gbits, c->d_gval, c->d_gmarg, c->d_gH, c->g_NW);
    else
        mi_pairs_kernel<<<refine_grid(c, npairs), EX_WARPS * 32, 0, c->stream>>>(c->d_colmajor, c->d_w, c->d_masks, d_pi, d_pj,
                                                                                npairs, c->n_pad, pc, d_mi, nullptr, nullptr, nullptr, 0);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_PAIRS1], c->stream));
    c->tm_pending |= TMP_PAIRS;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// Refines the candidate list d_in[0, n): class buckets (counting sort into d_bucket), one converged cooperative launch per class
// pair; d_refined / d_refined_wo run parallel to d_bucket. *n_valid: entries written (the int8 engine leaves invalid-marked
// holes in its list, dropped here). d_cnt: 32 u64 of scratch. Needs the grouped planes (g_state == 1).
int mi_refine_list(Ctx* c, const Candidate* d_in, uint64_t n, double pc, double wo_threshold, Candidate* d_bucket,
                   spk_edge* d_refined, double* d_refined_wo, unsigned long long* d_cnt, uint64_t* n_valid) {
    SPK_CUDA(c, cudaMemsetAsync(d_cnt, 0, 32 * sizeof(unsigned long long), c->stream));
    const unsigned cgrid = (unsigned)std::min<unsigned long long>((n + 2047) / 2048, (unsigned long long)c->sm_count * 8);
    classify_count_kernel<<<cgrid, 256, 0, c->stream>>>(d_in, n, c->d_masks, d_cnt);
    unsigned long long h_cnt[9], h_off[10];
    SPK_CUDA(c, cudaMemcpyAsync(h_cnt, d_cnt, sizeof h_cnt, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    h_off[0] = 0; for (int q = 0; q < 9; ++q) h_off[q + 1] = h_off[q] + h_cnt[q];
    if (n_valid) *n_valid = h_off[9];
    SPK_CUDA(c, cudaMemcpyAsync(d_cnt + 16, h_off, 9 * sizeof(unsigned long long), cudaMemcpyHostToDevice, c->stream));
    classify_scatter_kernel<<<cgrid, 256, 0, c->stream>>>(d_in, n, c->d_masks, d_cnt + 16, d_bucket);
    c->tm.kernel_launches += 2;
    for (int q = 0; q < 9; ++q) {
        const unsigned long long n_q = h_cnt[q], o_q = h_off[q];
        if (!n_q) continue;
        const unsigned g = (unsigned)((n_q + 127) / 128);
        static const bool coop = [] { const char* e = getenv("SPYDRPICK_B200_REFINE_COOP"); return !(e && e[0] == '0'); }();
#define SPK_RCASE(Q, A, B) case Q: \
            if (coop) mi_refine_coop_kernel<A, B, ((A) * (B) <= 4 ? 8 : ((A) * (B) == 8 ? 4 : 2))><<<g, 128, 0, c->stream>>>(c->d_masks, d_bucket + o_q, n_q, pc, c->w_total, wo_threshold, \
                d_refined + o_q, d_refined_wo + o_q, c->d_counters, c->d_gbits, c->d_gval, c->d_gmarg, c->d_gH, c->g_NW); \
            else mi_refine_class_kernel<A, B><<<g, 128, 0, c->stream>>>(c->d_masks, d_bucket + o_q, n_q, pc, c->w_total, wo_threshold, \
                d_refined + o_q, d_refined_wo + o_q, c->d_counters, c->d_gbits, c->d_gval, c->d_gmarg, c->d_gH, c->g_NW); \
            break;
        switch (q) {
            SPK_RCASE(0, 1, 1) SPK_RCASE(1, 1, 2) SPK_RCASE(2, 1, 4)
            SPK_RCASE(3, 2, 1) SPK_RCASE(4, 2, 2) SPK_RCASE(5, 2, 4)
            SPK_RCASE(6, 4, 1) SPK_RCASE(7, 4, 2) SPK_RCASE(8, 4, 4)
        }
#undef SPK_RCASE
        c->tm.kernel_launches += 1;
    }
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// the scan's own candidate list (c->d_cand -> c->d_refined, bucketed copy in c->d_cand2)
int mi_refine_candidates(Ctx* c, uint64_t ncand, double pc, double wo_threshold, uint64_t* n_valid) {
    if (n_valid) *n_valid = ncand;
    if (ncand == 0) return SPK_OK;
    { int rg = ensure_grouped(c, pc); if (rg) return rg; }
    if (c->g_state == 1) {
        unsigned long long* d_cnt = (unsigned long long*)scratch(c, 32 * sizeof(unsigned long long));
        if (!d_cnt) return fail(c, SPK_E_NOMEM, "scratch allocation failed");
        if (c->cand2_cap < c->cand_cap) {
            if (c->d_cand2) { cudaFree(c->d_cand2); c->d_cand2 = nullptr; c->cand2_cap = 0; }
            if (cudaMalloc(&c->d_cand2, c->cand_cap * sizeof(Candidate)) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "candidate sort buffer"); }
            c->cand2_cap = c->cand_cap;
        }
        return mi_refine_list(c, c->d_cand, ncand, pc, wo_threshold, c->d_cand2, c->d_refined, c->d_refined_wo, d_cnt, n_valid);
    }
    mi_refine_kernel<<<refine_grid(c, ncand), EX_WARPS * 32, 0, c->stream>>>(c->d_colmajor, c->d_w, c->d_masks, c->d_cand, ncand,
                                                                            c->n_pad, pc, c->d_refined, c->d_refined_wo,
                                                                            c->d_counters, nullptr, nullptr, nullptr, 0);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// MI of `total` pairs in `nseg` equal segments through the cooperative refine kernels: out[seg * per + k] = the MI values of
// segment seg in arbitrary order (the threshold estimator sorts every segment anyway). work = caller's scratch:
// 2 * total Candidates + total spk_edge + total doubles + 32 + nseg u64. Returns SPK_E_INVALID when the planes are not grouped.
__global__ void pairs_to_candidates_kernel(const uint32_t* __restrict__ pi, const uint32_t* __restrict__ pj, unsigned long long total,
                                           unsigned long long per, Candidate* __restrict__ out) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= total) return;
    Candidate cd; cd.i = pi[p]; cd.j = pj[p]; cd.mi32 = __int_as_float(0x7fc00000); cd.pad = (uint32_t)(p / per);
    out[p] = cd;
}
__global__ void scatter_segment_mi_kernel(const Candidate* __restrict__ bucket, const spk_edge* __restrict__ refined, unsigned long long n,
                                          unsigned long long per, unsigned long long* __restrict__ seg_fill, double* __restrict__ out) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = p < n;
    const uint32_t seg = live ? bucket[p].pad : 0xffffffffu;
    // one atomic per (warp, segment): the lanes of a segment are ranked inside their match group (a million single atomics on
    // ten addresses had cost 0.64 ms)
    const uint32_t lane = threadIdx.x % 32;
    const uint32_t peers = __match_any_sync(0xffffffffu, seg);
    const uint32_t leader = __ffs(peers) - 1u;
    unsigned long long base = 0;
    if (live && lane == leader) base = atomicAdd(&seg_fill[seg], (unsigned long long)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (live) {
        const unsigned long long k = base + __popc(peers & ((1u << lane) - 1u));
        if (k < per) out[(unsigned long long)seg * per + k] = refined[p].mi;
    }
}
size_t mi_pairs_segmented_scratch(uint64_t total, uint32_t nseg) {
    return (size_t)total * (2 * sizeof(Candidate) + sizeof(spk_edge) + 8) + (32 + (size_t)nseg) * 8 + 1024;
}
int mi_pairs_segmented(Ctx* c, const uint32_t* d_pi, const uint32_t* d_pj, uint64_t total, uint64_t per, uint32_t nseg, double pc,
                       double* d_mi, void* work) {
    if (total == 0) return SPK_OK;
    { int rg = ensure_grouped(c, pc); if (rg) return rg; }
    if (c->g_state != 1) return SPK_E_INVALID;
    uint8_t* w = (uint8_t*)work;
    Candidate* d_in = (Candidate*)w; w += (size_t)total * sizeof(Candidate);
    Candidate* d_bucket = (Candidate*)w; w += (size_t)total * sizeof(Candidate);
    spk_edge* d_ref = (spk_edge*)w; w += (size_t)total * sizeof(spk_edge);
    double* d_wo = (double*)w; w += (size_t)total * 8;
    w = (uint8_t*)(((uintptr_t)w + 255) & ~(uintptr_t)255);
    unsigned long long* d_cnt = (unsigned long long*)w; unsigned long long* d_fill = d_cnt + 32;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_PAIRS0], c->stream));
    pairs_to_candidates_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(d_pi, d_pj, total, per, d_in);
    SPK_CUDA(c, cudaMemsetAsync(d_fill, 0, (size_t)nseg * 8, c->stream));
    uint64_t nv = 0;
    int r = mi_refine_list(c, d_in, total, pc, INFINITY, d_bucket, d_ref, d_wo, d_cnt, &nv); if (r) return r;
    if (nv != total) return fail(c, SPK_E_INVALID, "mi_pairs_segmented: %llu of %llu pairs classified", (unsigned long long)nv, (unsigned long long)total);
    scatter_segment_mi_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(d_bucket, d_ref, total, per, d_fill, d_mi);
    c->tm.kernel_launches += 2;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_PAIRS1], c->stream));
    c->tm_pending |= TMP_PAIRS;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// audit_cand_lo: NaN = the list holds no audit entries; else thr - band of the filter (entries marked pad = 1 are judged, not saved)
int mi_compact(Ctx* c, uint64_t ncand, const spk_scan_params& p, uint64_t circle, float audit_cand_lo) {
    if (ncand == 0) return SPK_OK;
    const Candidate* meta = (audit_cand_lo == audit_cand_lo) ? (c->g_state == 1 ? c->d_cand2 : c->d_cand) : nullptr;
    mi_compact_kernel<<<(unsigned)((ncand + 255) / 256), 256, 0, c->stream>>>(
        c->d_refined, c->d_refined_wo, ncand, c->d_pos, p.threshold, p.ld_threshold, p.linear_genome, circle,
        (unsigned long long*)c->d_colmax, c->d_edges, c->edges_cap, c->d_wo, c->wo_cap, c->d_counters, meta, audit_cand_lo);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// exhaustive diagnostic engine: every pair of this partition goes through the fp64 kernels.
int mi_scan_exact(Ctx* c, const spk_scan_params& p, uint64_t circle) {
    const uint32_t L = c->L;
    const uint32_t nparts = p.n_parts ? p.n_parts : 1;
    // rows are dealt to partitions in contiguous chunks of 64 rows, round-robin
    const uint32_t chunk_rows = 64;
    float ms_refine = 0.f, ms_compact = 0.f;
    for (uint32_t r0 = 0, chunk = 0; r0 + 1 < L; r0 += chunk_rows, ++chunk) {
        if (chunk % nparts != p.part_index) continue;
        const uint32_t r1 = (r0 + chunk_rows < L - 1) ? r0 + chunk_rows : L - 1;
        // split so that the chunk fits the candidate buffer
        uint32_t a = r0;
        while (a < r1) {
            uint32_t b = a; unsigned long long cnt = 0;
            while (b < r1 && cnt + (L - 1 - b) <= c->cand_cap) { cnt += (L - 1 - b); ++b; }
            if (b == a) return fail(c, SPK_E_OVERFLOW, "candidate buffer smaller than one row of pairs");
            enumerate_pairs_kernel<<<b - a, 256, 0, c->stream>>>(c->d_cand, a, b, L, cnt);
            c->tm.kernel_launches += 1;
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_REF0], c->stream));
            int r = mi_refine_candidates(c, cnt, p.pseudocount, p.threshold); if (r) return r;
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_REF1], c->stream));
            r = mi_compact(c, cnt, p, circle, __builtin_nanf("")); if (r) return r;
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_CMP1], c->stream));
            SPK_CUDA(c, cudaEventSynchronize(c->ev[EV_CMP1]));
            float t1, t2;
            SPK_CUDA(c, cudaEventElapsedTime(&t1, c->ev[EV_REF0], c->ev[EV_REF1]));
            SPK_CUDA(c, cudaEventElapsedTime(&t2, c->ev[EV_REF1], c->ev[EV_CMP1]));
            ms_refine += t1; ms_compact += t2;
            a = b;
        }
    }
    c->tm.refine_ms = ms_refine; c->tm.compact_ms = ms_compact;
    return SPK_OK;
}

} // namespace spk
