// spk_mi_gemm.cu -- K2: all-pairs weighted joint counts as a tcgen05 (5th-gen tensor core) GEMM
// with a fused fp32 MI / threshold / colmax-candidate / ballot-compaction epilogue.
//
// Replaces the reference's hot loops
//   Weighted_crosstable_2Dblock::operator()   externals/apegrunt/include/misc/Crosstable.hpp:362-427
//   mutual_information_block_kernel::block     include/mi.hpp:108-181
//   MI_solver::operator() store loop           include/mi.hpp:414-461
// as a FILTER: every pair whose fp32 MI is within `band` of the save threshold or of the running
// per-column maximum is emitted as a candidate; candidates are re-evaluated in fp64 with the
// reference's exact formula by spk_mi_exact.cu, so reported MI values and the saved-edge set do
// not depend on tensor-core rounding.
//
// GEMM formulation (per output tile = 32 i-columns x 64 j-columns):
//   A (M=128)  rows 4*i+a : V_t[s] = [x_s,i == a] * w_t[s]  bf16, t < T terms of the weight split
//   B (N=256)  rows 4*j+b : U[s]   = [x_s,j == b]           bf16 (exact)
//   D[4i+a][4j+b] = sum_t sum_s V_t * U  = weighted joint count of (a,b), fp32 in TMEM.
// Gap row/column/corner of the 5x5 table come from the fp64-precomputed column marginals.
//
// Kernel anatomy (persistent, 1 CTA/SM, 192 threads):
//   warp 0    TMA producer      cp.async.bulk.tensor (128B swizzle) -> smem ring, mbarrier expect_tx
//   warp 1    MMA issuer        one elected lane: tcgen05.mma.cta_group::1.kind::f16 M128 N256 K16,
//                               tcgen05.commit -> smem "empty" / TMEM "full" barriers; owns TMEM alloc
//   warps 2-5 epilogue          tcgen05.ld 32x32b.x16, quad transpose (shfl), MI math, candidates
// TMEM: 2 accumulator buffers x 256 columns (all 512) so the epilogue of tile t overlaps the
// mainloop of tile t+1.
#include <cuda.h>
#include <algorithm>
#include <cmath>
#include <cstring>
#include "spk_internal.h"
#include "spk_tcgen05.cuh"

namespace spk {

constexpr int GEMM_THREADS = 192;
constexpr uint32_t A_TILE_BYTES = 128 * BLOCK_K * 2;      // 16 KB
constexpr uint32_t B_TILE_BYTES = 256 * BLOCK_K * 2;      // 32 KB
constexpr uint32_t UMMA_IDESC =                            // cute::UMMA::InstrDescriptor (mma_sm100_desc.hpp)
    (1u << 4)  |          // c_format  = F32
    (1u << 7)  |          // a_format  = BF16
    (1u << 10) |          // b_format  = BF16
    (0u << 15) | (0u << 16) |   // A, B K-major
    ((256u >> 3) << 17) | // N = 256
    ((128u >> 4) << 24);  // M = 128

struct GemmParams {
    const ColStat* colstat;
    Candidate* cand; unsigned long long cand_cap;
    unsigned long long* counters;      // [0] n_cand, [4] overflow flag
    uint32_t* colmax32;                // running per-column fp32 maxima (ordered-uint encoding)
    const unsigned long long* sr_prefix;   // [n_sr+1] chunk prefix per super-row
    uint32_t n_sr, n_i, n_j, L;
    uint32_t num_kb;                   // K blocks of 64 samples
    uint32_t rows_per_term;            // 4*L_pad (row offset between weight terms in V)
    uint32_t part, n_parts;
    unsigned long long n_local_slots;
    float w_total, pc, thr_lo, thr_hi, band;
    unsigned long long ld_threshold, circle; int linear;
    float* debug_out;                  // non-null: single tile (debug_ib, debug_jb), raw counts dump
    uint32_t debug_ib, debug_jb;
};

// ---------------------------------------------------------------------------------------------
// tile schedule: local slot -> (ib, jb). 16 consecutive slots share the j panel (L2 reuse).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ bool slot_to_tile(const GemmParams& p, unsigned long long ls, uint32_t& ib, uint32_t& jb) {
    if (p.debug_out) { ib = p.debug_ib; jb = p.debug_jb; return ls == 0; }
    const unsigned long long q = (ls / SUPER_I) * p.n_parts + p.part;
    const uint32_t w = (uint32_t)(ls % SUPER_I);
    if (q >= p.sr_prefix[p.n_sr]) return false;
    uint32_t lo = 0, hi = p.n_sr;                      // find sr with prefix[sr] <= q < prefix[sr+1]
    while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (p.sr_prefix[mid] <= q) lo = mid; else hi = mid; }
    ib = lo * SUPER_I + w;
    jb = lo * (SUPER_I / 2) + (uint32_t)(q - p.sr_prefix[lo]);
    return ib < p.n_i && jb < p.n_j && jb >= ib / 2;
}

// ---------------------------------------------------------------------------------------------
// fp32 MI of one pair from its 4x4 non-gap joint counts t[a][b] (a: state of column i).
// Same quantity as include/mi.hpp:146-181 with full presence masks:
//   MI = sum_{a in Pi, b in Pj} p_ab ln p_ab - sum_a p_a ln p_a - sum_b p_b ln p_b,
//   p_ab = (N_ab + pc)/Z, Z = W + pc|Pi||Pj|; the two marginal sums are precomputed per column.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float xlog2x(float p) { return p * __log2f(p); }

__device__ __forceinline__ float pair_mi32(const float* t, const float* mi_, uint32_t mask_i, float hi_k,
                                           const float* mj_, uint32_t mask_j, float hj_k, float pc, float invZ) {
    float colsum[4] = {0.f, 0.f, 0.f, 0.f};
    float acc = 0.f;
    float gapj_sum = 0.f;
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        float rowsum = 0.f, racc = 0.f;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const float nab = t[a * 4 + b];
            rowsum += nab; colsum[b] += nab;
            const float term = xlog2x((nab + pc) * invZ);
            racc += ((mask_j >> b) & 1u) ? term : 0.f;
        }
        const float ngap = fmaxf(mi_[a] - rowsum, 0.f);                 // cell (a, gap_j)
        const float tg = xlog2x((ngap + pc) * invZ);
        racc += ((mask_j >> 4) & 1u) ? tg : 0.f;
        acc += ((mask_i >> a) & 1u) ? racc : 0.f;
    }
    if ((mask_i >> 4) & 1u) {                                           // gap_i row
        float racc = 0.f;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const float ngap = fmaxf(mj_[b] - colsum[b], 0.f);          // cell (gap_i, b)
            gapj_sum += ngap;
            const float term = xlog2x((ngap + pc) * invZ);
            racc += ((mask_j >> b) & 1u) ? term : 0.f;
        }
        const float corner = fmaxf(mi_[4] - gapj_sum, 0.f);             // cell (gap_i, gap_j)
        const float tc = xlog2x((corner + pc) * invZ);
        racc += ((mask_j >> 4) & 1u) ? tc : 0.f;
        acc += racc;
    }
    return acc * 0.69314718055994530942f - hi_k - hj_k;
}

// ---------------------------------------------------------------------------------------------
template <int T, int NSTAGES>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
mi_gemm_kernel(const __grid_constant__ CUtensorMap map_v, const __grid_constant__ CUtensorMap map_u, const GemmParams p) {
    constexpr uint32_t STAGE_BYTES = B_TILE_BYTES + T * A_TILE_BYTES;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // align inside the shared window with pointer arithmetic on the __shared__ symbol (keeps LDS/STS addressing)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* tiles = smem;                                       // NSTAGES * STAGE_BYTES
    uint8_t* aux = smem + (size_t)NSTAGES * STAGE_BYTES;
    ColStat* sI = (ColStat*)aux;                                  // [32]
    ColStat* sJ = sI + TILE_I;                                    // [64]
    uint32_t* s_cmj = (uint32_t*)(sJ + TILE_J);                   // [64] running colmax of the tile's j columns
    float* s_mi = (float*)(s_cmj + TILE_J);                       // [16][128] fp32 MI of the tile (pass 1 -> pass 2)
    uint64_t* full_bar = (uint64_t*)(s_mi + 16 * 128);
    uint64_t* empty_bar = full_bar + NSTAGES;
    uint64_t* tfull_bar = empty_bar + NSTAGES;                    // [2]
    uint64_t* tempty_bar = tfull_bar + 2;                         // [2]
    uint32_t* tmem_slot = (uint32_t*)(tempty_bar + 2);

    const uint32_t warp = threadIdx.x / 32, lane = threadIdx.x % 32;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NSTAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tfull_bar[b], 1); mbar_init(&tempty_bar[b], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_u) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t stage = 0, phase = 0;
            for (unsigned long long ls = blockIdx.x; ls < p.n_local_slots; ls += gridDim.x) {
                uint32_t ib, jb;
                if (!slot_to_tile(p, ls, ib, jb)) continue;
                for (uint32_t kb = 0; kb < p.num_kb; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1u);
                    uint8_t* st = tiles + (size_t)stage * STAGE_BYTES;
                    mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
                    tma_load_2d(&map_u, &full_bar[stage], st, (int32_t)(kb * BLOCK_K), (int32_t)(jb * 256u));
#pragma unroll
                    for (int t = 0; t < T; ++t)
                        tma_load_2d(&map_v, &full_bar[stage], st + B_TILE_BYTES + t * A_TILE_BYTES, (int32_t)(kb * BLOCK_K),
                                    (int32_t)(t * p.rows_per_term + ib * 128u));
                    if (++stage == NSTAGES) { stage = 0; phase ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            uint32_t stage = 0, phase = 0, it = 0;
            for (unsigned long long ls = blockIdx.x; ls < p.n_local_slots; ls += gridDim.x) {
                uint32_t ib, jb;
                if (!slot_to_tile(p, ls, ib, jb)) continue;
                const uint32_t buf = it & 1u, acc_phase = (it >> 1) & 1u;
                mbar_wait(&tempty_bar[buf], acc_phase ^ 1u);
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + buf * 256u;
                for (uint32_t kb = 0; kb < p.num_kb; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t sb = smem_u32(tiles + (size_t)stage * STAGE_BYTES);
#pragma unroll
                    for (int t = 0; t < T; ++t) {
                        const uint32_t sa = sb + B_TILE_BYTES + t * A_TILE_BYTES;
#pragma unroll
                        for (int k = 0; k < BLOCK_K / 16; ++k) {
                            const uint64_t adesc = umma_smem_desc(sa + k * 32);
                            const uint64_t bdesc = umma_smem_desc(sb + k * 32);
                            umma_bf16(d_tmem, adesc, bdesc, UMMA_IDESC, (kb | (uint32_t)t | (uint32_t)k) ? 1u : 0u);
                        }
                    }
                    umma_commit(&empty_bar[stage]);              // frees the smem stage when these MMAs retire
                    if (++stage == NSTAGES) { stage = 0; phase ^= 1u; }
                }
                umma_commit(&tfull_bar[buf]);                    // accumulator complete -> epilogue
                ++it;
            }
        }
    } else {
        // ===================== epilogue (4 warps, 128 threads) =====================
        const uint32_t q = warp & 3u;                            // TMEM lane quadrant this warp may access
        const uint32_t et = (warp - 2) * 32 + lane;              // 0..127 epilogue thread id
        const uint32_t i_local = q * 8 + lane / 4;
        const uint32_t q4 = lane & 3u;
        uint32_t it = 0;
        for (unsigned long long ls = blockIdx.x; ls < p.n_local_slots; ls += gridDim.x) {
            uint32_t ib, jb;
            if (!slot_to_tile(p, ls, ib, jb)) continue;
            const uint32_t buf = it & 1u, acc_phase = (it >> 1) & 1u;
            // stage column statistics of this tile (previous tile's readers are past the trailing barrier)
            {
                const uint4* src_i = (const uint4*)(p.colstat + (size_t)ib * TILE_I);
                const uint4* src_j = (const uint4*)(p.colstat + (size_t)jb * TILE_J);
                uint4* dst_i = (uint4*)sI; uint4* dst_j = (uint4*)sJ;
                dst_i[et] = src_i[et];                           // 32*64/16 = 128 uint4
                dst_j[et] = src_j[et]; dst_j[et + 128] = src_j[et + 128];
                if (et < TILE_J) s_cmj[et] = p.colmax32[jb * TILE_J + et];
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            const uint32_t gi = ib * TILE_I + i_local;
            float m_i[5];
#pragma unroll
            for (int a = 0; a < 5; ++a) m_i[a] = sI[i_local].m[a];
            const uint32_t mask_i = sI[i_local].mask, np_i = sI[i_local].npres;
            const unsigned long long pos_i = ((unsigned long long)sI[i_local].pos_hi << 32) | sI[i_local].pos_lo;
            float cm_i = f32_from_ordered(p.colmax32[gi]);
            float cm_i_new = cm_i;
            uint32_t ldmask = 0;

            mbar_wait(&tfull_bar[buf], acc_phase);
            tcgen05_fence_after();
            const uint32_t taddr0 = tmem_base + buf * 256u + ((q * 32u) << 16);
#pragma unroll 1
            for (uint32_t c = 0; c < 16; ++c) {
                float v[16];
                tmem_ld16(taddr0 + c * 16u, v);
                // quad transpose: lane (a = lane%4) holds v[jj][b]; afterwards v[a'][b] of j = 4c + lane%4
                {
                    const bool b0 = lane & 1u, b1 = lane & 2u;
#pragma unroll
                    for (int pr = 0; pr < 2; ++pr)
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            const int s0 = (2 * pr) * 4 + b, s1 = (2 * pr + 1) * 4 + b;
                            const float send = b0 ? v[s0] : v[s1];
                            const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
                            if (b0) v[s0] = recv; else v[s1] = recv;
                        }
#pragma unroll
                    for (int pr = 0; pr < 2; ++pr)
#pragma unroll
                        for (int b = 0; b < 4; ++b) {
                            const int s0 = pr * 4 + b, s1 = (pr + 2) * 4 + b;
                            const float send = b1 ? v[s0] : v[s1];
                            const float recv = __shfl_xor_sync(0xffffffffu, send, 2);
                            if (b1) v[s0] = recv; else v[s1] = recv;
                        }
                }
                const uint32_t j_local = c * 4 + q4;
                const uint32_t gj = jb * TILE_J + j_local;
                if (p.debug_out) {
#pragma unroll
                    for (int k = 0; k < 16; ++k) p.debug_out[((size_t)i_local * TILE_J + j_local) * 16 + k] = v[k];
                    continue;
                }
                const ColStat& cj = sJ[j_local];
                const uint32_t mask_j = cj.mask, np_j = cj.npres;
                const bool valid = (gi < gj) && (gj < p.L) && np_i && np_j;
                float mi32 = __int_as_float(0x7fc00000);                   // NaN marks "no pair here"
                if (valid) {
                    float m_j[4] = {cj.m[0], cj.m[1], cj.m[2], cj.m[3]};
                    const float invZ = __frcp_rn(p.w_total + p.pc * (float)(np_i * np_j));
                    mi32 = pair_mi32(v, m_i, mask_i, sI[i_local].h[np_j - 1], m_j, mask_j, cj.h[np_i - 1], p.pc, invZ);
                    const unsigned long long pos_j = ((unsigned long long)cj.pos_hi << 32) | cj.pos_lo;
                    if (genome_distance(pos_i, pos_j, p.linear, p.circle) > p.ld_threshold) {
                        ldmask |= 1u << c;
                        cm_i_new = fmaxf(cm_i_new, mi32);
                        if (mi32 > f32_from_ordered(s_cmj[j_local])) atomicMax(&s_cmj[j_local], f32_ordered(mi32));
                    }
                }
                s_mi[c * 128 + et] = mi32;
            }
            // accumulator drained: hand the TMEM buffer back to the MMA warp
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[buf]);
            if (!p.debug_out) {
                // tile-level maxima of this thread's i column (4 lanes of the quad cover different j)
                cm_i_new = fmaxf(cm_i_new, __shfl_xor_sync(0xffffffffu, cm_i_new, 1));
                cm_i_new = fmaxf(cm_i_new, __shfl_xor_sync(0xffffffffu, cm_i_new, 2));
                if (q4 == 0 && cm_i_new > cm_i) atomicMax(&p.colmax32[gi], f32_ordered(cm_i_new));
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");                 // s_cmj now holds the tile-level j maxima
            if (!p.debug_out) {
                if (et < TILE_J) atomicMax(&p.colmax32[jb * TILE_J + et], s_cmj[et]);
                // pass 2: candidate emission. A pair is a candidate if its fp32 MI is within `band` of the save
                // threshold, or (LD-admissible) within `band` of the running maximum of one of its columns while
                // that maximum has not yet cleared thr+band (columns above it are covered by the threshold rule).
                const bool i_open = !(cm_i_new > p.thr_hi);
#pragma unroll 1
                for (uint32_t c = 0; c < 16; ++c) {
                    const float mi32 = s_mi[c * 128 + et];
                    const uint32_t j_local = c * 4 + q4;
                    bool emit = mi32 > p.thr_lo;                           // false for NaN
                    if ((ldmask >> c) & 1u) {
                        const float cm_j = f32_from_ordered(s_cmj[j_local]);
                        emit = emit || (i_open && mi32 >= cm_i_new - p.band) || (!(cm_j > p.thr_hi) && mi32 >= cm_j - p.band);
                    }
                    const uint32_t bal = __ballot_sync(0xffffffffu, emit);
                    if (bal) {
                        unsigned long long base = 0;
                        if (lane == 0) base = atomicAdd(&p.counters[0], (unsigned long long)__popc(bal));
                        base = __shfl_sync(0xffffffffu, base, 0);
                        if (emit) {
                            const unsigned long long idx = base + __popc(bal & ((1u << lane) - 1u));
                            if (idx < p.cand_cap) { Candidate cd; cd.i = gi; cd.j = jb * TILE_J + j_local; cd.mi32 = mi32; cd.pad = 0; p.cand[idx] = cd; }
                            else p.counters[4] = 1ull;
                        }
                    }
                }
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");                 // sI/sJ/s_cmj/s_mi free for the next tile
            ++it;
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// =============================================================================================
// host side
// =============================================================================================
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)ptr;
    }
    return fn;
}

// 2D bf16 tensor [rows][k_elems] (k contiguous), box = 64 k-elements (128 B) x box_rows, 128B swizzle
static int make_map(Ctx* c, CUtensorMap* map, void* base, uint64_t rows, uint64_t k_elems, uint32_t box_rows) {
    PFN_encodeTiled enc = get_encode_fn();
    if (!enc) return fail(c, SPK_E_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t gdim[2] = {k_elems, rows};
    cuuint64_t gstride[1] = {k_elems * 2};
    cuuint32_t box[2] = {(cuuint32_t)BLOCK_K, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(c, SPK_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
    return SPK_OK;
}

__global__ void fill_u32_kernel(uint32_t* p, uint32_t v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

template <int T, int NSTAGES>
static int launch_gemm(Ctx* c, const CUtensorMap& mv, const CUtensorMap& mu, const GemmParams& gp, unsigned grid) {
    constexpr size_t STAGE_BYTES = B_TILE_BYTES + T * A_TILE_BYTES;
    const size_t smem = (size_t)NSTAGES * STAGE_BYTES + (TILE_I + TILE_J) * sizeof(ColStat) + TILE_J * 4 + 16 * 128 * 4 +
                        (2 * NSTAGES + 4) * 8 + 16 + 1024;
    auto kfn = mi_gemm_kernel<T, NSTAGES>;
    SPK_CUDA(c, cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kfn<<<grid, GEMM_THREADS, smem, c->stream>>>(mv, mu, gp);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

static int run_gemm(Ctx* c, GemmParams& gp, int T) {
    CUtensorMap mv, mu;
    int r = make_map(c, &mv, c->d_V, (uint64_t)T * 4 * c->L_pad, c->n_pad, 128); if (r) return r;
    r = make_map(c, &mu, c->d_U, (uint64_t)4 * c->L_pad, c->n_pad, 256); if (r) return r;
    gp.num_kb = c->n_pad / BLOCK_K;
    gp.rows_per_term = 4 * c->L_pad;
    gp.colstat = c->d_colstat;
    gp.L = c->L;
    unsigned long long slots = gp.debug_out ? 1ull : gp.n_local_slots;
    unsigned grid = (unsigned)(slots < (unsigned long long)c->sm_count ? (slots ? slots : 1) : (unsigned long long)c->sm_count);
    switch (T) {
        case 1: return launch_gemm<1, 4>(c, mv, mu, gp, grid);
        case 2: return launch_gemm<2, 3>(c, mv, mu, gp, grid);
        case 3: return launch_gemm<3, 2>(c, mv, mu, gp, grid);
        default: return fail(c, SPK_E_INVALID, "n_terms must be 1..3");
    }
}

int mi_gemm_filter(Ctx* c, const spk_scan_params& p, int T, double band, uint64_t circle) {
    GemmParams gp; memset(&gp, 0, sizeof gp);
    const uint32_t n_i = (c->L + TILE_I - 1) / TILE_I, n_j = c->L_pad / TILE_J;
    const uint32_t n_sr = (n_i + SUPER_I - 1) / SUPER_I;
    std::vector<unsigned long long> prefix(n_sr + 1, 0);
    for (uint32_t sr = 0; sr < n_sr; ++sr) {
        const uint32_t j0 = sr * (SUPER_I / 2);
        prefix[sr + 1] = prefix[sr] + (j0 < n_j ? n_j - j0 : 0);
    }
    const uint32_t nparts = p.n_parts ? p.n_parts : 1;
    if (p.part_index >= nparts) return fail(c, SPK_E_INVALID, "part_index >= n_parts");
    const unsigned long long total_chunks = prefix[n_sr];
    const unsigned long long local_chunks = total_chunks > p.part_index ? (total_chunks - p.part_index + nparts - 1) / nparts : 0;
    unsigned long long* d_prefix = (unsigned long long*)c->d_tilemap;
    SPK_CUDA(c, cudaMemcpyAsync(d_prefix, prefix.data(), (n_sr + 1) * 8, cudaMemcpyHostToDevice, c->stream));
    fill_u32_kernel<<<(c->L_pad + 255) / 256, 256, 0, c->stream>>>((uint32_t*)c->d_colmax32, f32_ordered(-INFINITY), c->L_pad);
    c->tm.kernel_launches += 1;
    gp.cand = c->d_cand; gp.cand_cap = c->cand_cap; gp.counters = c->d_counters;
    gp.colmax32 = (uint32_t*)c->d_colmax32;
    gp.sr_prefix = d_prefix; gp.n_sr = n_sr; gp.n_i = n_i; gp.n_j = n_j;
    gp.part = p.part_index; gp.n_parts = nparts; gp.n_local_slots = local_chunks * SUPER_I;
    gp.w_total = (float)c->w_total; gp.pc = (float)p.pseudocount;
    gp.band = (float)band; gp.thr_lo = (float)(p.threshold - band); gp.thr_hi = (float)(p.threshold + band);
    gp.ld_threshold = p.ld_threshold; gp.circle = circle; gp.linear = p.linear_genome;
    gp.debug_out = nullptr;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_GEMM0], c->stream));
    int r = run_gemm(c, gp, T); if (r) return r;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_GEMM1], c->stream));
    c->tm_pending |= TMP_GEMM;
    // tiles actually computed by this partition (valid slots)
    unsigned long long tiles = 0;
    for (unsigned long long lq = 0; lq < local_chunks; ++lq) {
        const unsigned long long q = lq * nparts + p.part_index;
        uint32_t sr = (uint32_t)(std::upper_bound(prefix.begin(), prefix.end(), q) - prefix.begin()) - 1;
        const uint32_t jb = sr * (SUPER_I / 2) + (uint32_t)(q - prefix[sr]);
        for (uint32_t w = 0; w < SUPER_I; ++w) { const uint32_t ib = sr * SUPER_I + w; if (ib < n_i && jb < n_j && jb >= ib / 2) ++tiles; }
    }
    c->tm.gemm_tiles = tiles; c->tm.gemm_pairs = tiles * (unsigned long long)(TILE_I * TILE_J);
    return SPK_OK;
}

int mi_gemm_debug_tile(Ctx* c, uint32_t ib, uint32_t jb, int T, float* d_out) {
    GemmParams gp; memset(&gp, 0, sizeof gp);
    gp.counters = c->d_counters; gp.colmax32 = (uint32_t*)c->d_colmax32;
    gp.sr_prefix = (const unsigned long long*)c->d_tilemap;
    gp.n_parts = 1; gp.n_local_slots = 1;
    gp.w_total = (float)c->w_total; gp.pc = 0.5f;
    gp.debug_out = d_out; gp.debug_ib = ib; gp.debug_jb = jb;
    gp.n_i = (c->L + TILE_I - 1) / TILE_I; gp.n_j = c->L_pad / TILE_J;
    int r = run_gemm(c, gp, T); if (r) return r;
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}

} // namespace spk
