// spk_hamming.cu -- K1: sample-pair identity counts on 3 bit planes (XOR / OR / POPC).
//
// Replaces apegrunt::calculate_weights' hot loop
//   (externals/apegrunt/include/apegrunt/Alignment_StateVector_weights.hpp:242-264) ->
//   is_similar_to_impl (.../StateVector_impl_block_compressed_alignment_storage.hpp:607-660) ->
//   count_identical (.../State_block.hpp:347-430).
// ident(i,j) = L - popc( (p0_i^p0_j) | (p1_i^p1_j) | (p2_i^p2_j) ) summed over ceil(L/32) words
// (pad bits are 0 in every plane, so they never count as mismatches).
// Integer arithmetic, bit-exact. Each block owns a 64x64 tile of sample pairs and streams the
// bit planes through shared memory in 16-word chunks; each thread keeps a 4x4 register tile.
//
// hamming_csa_kernel (default) balances the two pipes the loop needs (LOP3 on the ALU pipe, POPC on the quarter-rate
// XU pipe): in 3 of the 16 pair cells of a thread the mismatch words of two consecutive plane words go through a
// carry-save adder (ones' = ones ^ m1 ^ m2, carry = maj(ones, m1, m2); mismatches = 2 * sum popc(carry) + popc(ones)),
// i.e. ONE popcount per two words for one more LOP3; the other 13 cells count every word. The chunk is stored transposed and XOR-swizzled in shared memory ([plane][word][64 samples]),
// so a thread's four rows / four columns of one word are ONE 16-byte LDS (6 LDS.128 per word instead of 24 LDS.32),
// and chunks arrive by cp.async (zero-fill outside the alignment) into a two-stage ring. Only the upper-triangle
// tile pairs are launched. hamming_tile_kernel is the first version, kept for A/B runs (SPYDRPICK_B200_K1=0).
#include <cstdlib>
#include "spk_internal.h"

namespace spk {

constexpr int HT = 64;        // samples per tile side
constexpr int HCH = 16;       // words per chunk
constexpr int HPAD = HCH + 1; // bank-conflict-free row pitch

__global__ void __launch_bounds__(256)
hamming_tile_kernel(const uint32_t* __restrict__ planes, uint32_t n, uint32_t L, uint32_t W32,
                    unsigned long long min_identical, int quirk, unsigned long long* __restrict__ counts,
                    uint32_t part, uint32_t n_parts) {
    const uint32_t ti = blockIdx.y, tj = blockIdx.x;
    if (tj < ti) return;
    if (n_parts > 1 && (ti * gridDim.x + tj) % n_parts != part) return;        // tile pairs dealt round-robin to the partitions
    __shared__ uint32_t sI[3][HT][HPAD];
    __shared__ uint32_t sJ[3][HT][HPAD];
    __shared__ unsigned int cntI[HT], cntJ[HT];
    const uint32_t tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    const uint32_t i0 = ti * HT, j0 = tj * HT;
    const size_t plane = (size_t)n * W32;
    if (threadIdx.x < HT) { cntI[threadIdx.x] = 0; cntJ[threadIdx.x] = 0; }

    uint32_t acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[r][q] = 0;

    for (uint32_t w0 = 0; w0 < W32; w0 += HCH) {
        __syncthreads();
        for (uint32_t e = threadIdx.x; e < 3 * HT * HCH; e += 256) {
            const uint32_t k = e % HCH, r = (e / HCH) % HT, p = e / (HCH * HT);
            const uint32_t w = w0 + k;
            const uint32_t si = i0 + r, sj = j0 + r;
            sI[p][r][k] = (si < n && w < W32) ? planes[p * plane + (size_t)si * W32 + w] : 0u;
            sJ[p][r][k] = (sj < n && w < W32) ? planes[p * plane + (size_t)sj * W32 + w] : 0u;
        }
        __syncthreads();
#pragma unroll 4
        for (int k = 0; k < HCH; ++k) {
            uint32_t a0[4], a1[4], a2[4], b0[4], b1[4], b2[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) { a0[r] = sI[0][ty + 16 * r][k]; a1[r] = sI[1][ty + 16 * r][k]; a2[r] = sI[2][ty + 16 * r][k]; }
#pragma unroll
            for (int q = 0; q < 4; ++q) { b0[q] = sJ[0][tx + 16 * q][k]; b1[q] = sJ[1][tx + 16 * q][k]; b2[q] = sJ[2][tx + 16 * q][k]; }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    acc[r][q] += __popc((a0[r] ^ b0[q]) | (a1[r] ^ b1[q]) | (a2[r] ^ b2[q]));
        }
    }

    const bool do_quirk = quirk && (L % 32u == 0u) && W32 > 0;
    unsigned int myI[4] = {0, 0, 0, 0}, myJ[4] = {0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t si = i0 + ty + 16 * r;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t sj = j0 + tx + 16 * q;
            if (si < n && sj < n && si < sj) {
                unsigned long long eff = (unsigned long long)L - acc[r][q];
                if (do_quirk) {
                    // reference double-counts the final 32-column block when L % 32 == 0
                    const size_t oi = (size_t)si * W32 + (W32 - 1), oj = (size_t)sj * W32 + (W32 - 1);
                    const uint32_t d = (planes[oi] ^ planes[oj]) | (planes[plane + oi] ^ planes[plane + oj]) |
                                       (planes[2 * plane + oi] ^ planes[2 * plane + oj]);
                    eff += 32u - __popc(d);
                }
                if (eff >= min_identical) { myI[r]++; myJ[q]++; }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) if (myI[r]) atomicAdd(&cntI[ty + 16 * r], myI[r]);
#pragma unroll
    for (int q = 0; q < 4; ++q) if (myJ[q]) atomicAdd(&cntJ[tx + 16 * q], myJ[q]);
    __syncthreads();
    if (threadIdx.x < HT) {
        const uint32_t si = i0 + threadIdx.x, sj = j0 + threadIdx.x;
        if (si < n && cntI[threadIdx.x]) atomicAdd(&counts[si], (unsigned long long)cntI[threadIdx.x]);
        if (sj < n && cntJ[threadIdx.x]) atomicAdd(&counts[sj], (unsigned long long)cntJ[threadIdx.x]);
    }
}

// ---------------------------------------------------------------------------------------------
// CSA version. Shared layout per stage: [which: I/J][plane 0..2][word k 0..15][group g' 0..15][4 samples],
// sample r of the tile = 4*g + s is stored in group g' = g ^ k (the swizzle keeps the transposing cp.async
// stores at most 2-way bank-conflicted; the 16-byte reads are conflict-free / broadcast).
constexpr int HC_STAGE_WORDS = 2 * 3 * HCH * HT;            // 6144 words = 24 KB per stage
constexpr int HC_STAGES = 2;
constexpr int HC_SMEM_BYTES = HC_STAGES * HC_STAGE_WORDS * 4;

__device__ __forceinline__ void hc_cp_async4(uint32_t dst_smem, const void* src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst_smem), "l"(src), "r"(src_bytes) : "memory");
}

// Per-thread constants of the chunk copy. Thread t copies word k = t % 16 of rows r0 + 16*ri (r0 = t / 16, ri = 0..3) of
// the six (I/J, plane) sub-arrays: 24 cp.async of 4 bytes per chunk, lanes running over the 16 words of a row (64-byte
// runs). Everything that does not depend on the chunk is computed once, so that a copy costs a handful of instructions
// (the generic index arithmetic had cost more ALU work than the 48 LOP3 of the mismatch words it feeds).
// Pipe balance: LOP3 / IADD3 issue at 16 lanes per clock and SM sub-partition, POPC at 4. A plain cell costs 3 LOP3 + 1 POPC
// (+ half an IADD3) per word, a carry-save cell 4 LOP3 + half a POPC: 13 plain + 3 carry-save cells of the 4x4 register tile
// load the two pipes about equally (ncu of the all-carry-save version: ALU pipe 87 % busy, XU 29 %).
__host__ __device__ constexpr bool hc_csa_cell(int r, int q) { return q == 3 && r < 3; }

struct HcCopy {
    uint32_t offI, offJ;        // word offset of (row i0 + r0, word k) / (row j0 + r0, word k) inside plane 0
    uint32_t rstride;           // 16 rows further: 16 * W32 words
    uint32_t plane32;           // words per plane
    int limI, limJ;             // rows r0 + 16*ri exist while 16*ri < lim
    uint32_t dthread;           // byte offset of (k, low swizzle bits, r0 & 3) inside a (I/J, plane) sub-array
    uint32_t kh64;              // 64 * (k >> 2): xor-ed with 64 * ri (the high swizzle bits)
    uint32_t k;
};

__device__ __forceinline__ void hc_issue_chunk(uint32_t stage_base, const uint32_t* __restrict__ planes, const HcCopy& cp,
                                               uint32_t W32, uint32_t w0) {
    const bool wok = w0 + cp.k < W32;
#pragma unroll
    for (int it = 0; it < 24; ++it) {
        const int ri = it % 4, pw = it / 4, p = pw % 3, which = pw / 3;      // pw = which * 3 + plane
        const bool ok = wok && (16 * ri < (which ? cp.limJ : cp.limI));
        const uint32_t off = (which ? cp.offJ : cp.offI) + (uint32_t)ri * cp.rstride + (uint32_t)p * cp.plane32 + w0;
        const uint32_t dst = stage_base + cp.dthread + (uint32_t)pw * (HCH * HT * 4) + (cp.kh64 ^ (uint32_t)(64 * ri));
        hc_cp_async4(dst, planes + (ok ? off : 0u), ok ? 4u : 0u);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
}

__global__ void __launch_bounds__(256, 2)
hamming_csa_kernel(const uint32_t* __restrict__ planes, uint32_t n, uint32_t L, uint32_t W32, uint32_t nt,
                   unsigned long long min_identical, int quirk, unsigned long long* __restrict__ counts,
                   uint32_t part, uint32_t n_parts) {
    extern __shared__ __align__(16) uint32_t hc_smem[];
    __shared__ unsigned int cntI[HT], cntJ[HT];
    // linear index over the upper-triangle tile pairs (ti <= tj), dealt round-robin to the partitions
    const uint32_t b = part + blockIdx.x * n_parts;
    uint32_t ti = (uint32_t)(((float)(2 * nt + 1) - sqrtf(fmaxf((float)(2 * nt + 1) * (float)(2 * nt + 1) - 8.f * (float)b, 0.f))) * 0.5f);
    if (ti >= nt) ti = nt - 1;
    // row ti starts at ti*nt - ti(ti-1)/2
    while (ti > 0 && ti * nt - ti * (ti - 1) / 2 > b) --ti;
    while ((ti + 1) * nt - (ti + 1) * ti / 2 <= b) ++ti;
    const uint32_t tj = ti + (b - (ti * nt - ti * (ti - 1) / 2));
    const uint32_t tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    const uint32_t i0 = ti * HT, j0 = tj * HT;
    const size_t plane = (size_t)n * W32;
    if (threadIdx.x < HT) { cntI[threadIdx.x] = 0; cntJ[threadIdx.x] = 0; }
    const uint32_t smem_base = (uint32_t)__cvta_generic_to_shared(hc_smem);

    uint32_t acc2[4][4], ones[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) { acc2[r][q] = 0; ones[r][q] = 0; }

    const uint32_t nchunks = (W32 + HCH - 1) / HCH;
    HcCopy cp;
    {
        const uint32_t k = threadIdx.x % HCH, r0 = threadIdx.x / HCH;       // r0 = 0..15
        cp.k = k;
        cp.offI = (i0 + r0) * W32 + k; cp.offJ = (j0 + r0) * W32 + k;
        cp.rstride = 16u * W32; cp.plane32 = n * W32;
        cp.limI = (int)n - (int)(i0 + r0); cp.limJ = (int)n - (int)(j0 + r0);
        // row r = r0 + 16*ri: group r >> 2 = (r0 >> 2) + 4*ri, swizzled with k: low bits (r0 >> 2) ^ (k & 3), high bits ri ^ (k >> 2)
        cp.dthread = 4u * (k * HT + 4u * ((r0 >> 2) ^ (k & 3u)) + (r0 & 3u));
        cp.kh64 = 64u * (k >> 2);
    }
    hc_issue_chunk(smem_base, planes, cp, W32, 0);
    for (uint32_t ch = 0; ch < nchunks; ++ch) {
        if (ch + 1 < nchunks) {
            hc_issue_chunk(smem_base + ((ch + 1) % HC_STAGES) * HC_STAGE_WORDS * 4, planes, cp, W32, (ch + 1) * HCH);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        const uint4* st = reinterpret_cast<const uint4*>(hc_smem + (ch % HC_STAGES) * HC_STAGE_WORDS);
        // uint4 index of (which, plane, k, group g) = ((which*3 + plane) * HCH + k) * 16 + (g ^ k)
#pragma unroll 2
        for (int k = 0; k < HCH; k += 2) {
            uint32_t m1[4][4];
            {
                const uint4 a0 = st[((0 * HCH) + k) * 16 + (ty ^ k)], a1 = st[((1 * HCH) + k) * 16 + (ty ^ k)], a2 = st[((2 * HCH) + k) * 16 + (ty ^ k)];
                const uint4 b0 = st[((3 * HCH) + k) * 16 + (tx ^ k)], b1 = st[((4 * HCH) + k) * 16 + (tx ^ k)], b2 = st[((5 * HCH) + k) * 16 + (tx ^ k)];
                const uint32_t A0[4] = {a0.x, a0.y, a0.z, a0.w}, A1[4] = {a1.x, a1.y, a1.z, a1.w}, A2[4] = {a2.x, a2.y, a2.z, a2.w};
                const uint32_t B0[4] = {b0.x, b0.y, b0.z, b0.w}, B1[4] = {b1.x, b1.y, b1.z, b1.w}, B2[4] = {b2.x, b2.y, b2.z, b2.w};
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int q = 0; q < 4; ++q) m1[r][q] = (A0[r] ^ B0[q]) | (A1[r] ^ B1[q]) | (A2[r] ^ B2[q]);
            }
            {
                const int k1 = k + 1;
                const uint4 a0 = st[((0 * HCH) + k1) * 16 + (ty ^ k1)], a1 = st[((1 * HCH) + k1) * 16 + (ty ^ k1)], a2 = st[((2 * HCH) + k1) * 16 + (ty ^ k1)];
                const uint4 b0 = st[((3 * HCH) + k1) * 16 + (tx ^ k1)], b1 = st[((4 * HCH) + k1) * 16 + (tx ^ k1)], b2 = st[((5 * HCH) + k1) * 16 + (tx ^ k1)];
                const uint32_t A0[4] = {a0.x, a0.y, a0.z, a0.w}, A1[4] = {a1.x, a1.y, a1.z, a1.w}, A2[4] = {a2.x, a2.y, a2.z, a2.w};
                const uint32_t B0[4] = {b0.x, b0.y, b0.z, b0.w}, B1[4] = {b1.x, b1.y, b1.z, b1.w}, B2[4] = {b2.x, b2.y, b2.z, b2.w};
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint32_t m2 = (A0[r] ^ B0[q]) | (A1[r] ^ B1[q]) | (A2[r] ^ B2[q]);
                        const uint32_t x = m1[r][q];
                        if (hc_csa_cell(r, q)) {
                            const uint32_t o = ones[r][q];
                            acc2[r][q] += __popc((o & x) | (o & m2) | (x & m2));    // carry: weight 2
                            ones[r][q] = o ^ x ^ m2;
                        } else {
                            acc2[r][q] += __popc(x) + __popc(m2);                    // plain cell: one IADD3 for two popcounts
                        }
                    }
            }
        }
        __syncthreads();                                  // the stage is refilled by the next iteration's cp.async
    }

    const bool do_quirk = quirk && (L % 32u == 0u) && W32 > 0;
    unsigned int myI[4] = {0, 0, 0, 0}, myJ[4] = {0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t si = i0 + 4 * ty + r;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t sj = j0 + 4 * tx + q;
            if (si < n && sj < n && si < sj) {
                unsigned long long eff = (unsigned long long)L - (hc_csa_cell(r, q) ? 2u * acc2[r][q] + (uint32_t)__popc(ones[r][q]) : acc2[r][q]);
                if (do_quirk) {
                    // reference double-counts the final 32-column block when L % 32 == 0
                    const size_t oi = (size_t)si * W32 + (W32 - 1), oj = (size_t)sj * W32 + (W32 - 1);
                    const uint32_t d = (planes[oi] ^ planes[oj]) | (planes[plane + oi] ^ planes[plane + oj]) |
                                       (planes[2 * plane + oi] ^ planes[2 * plane + oj]);
                    eff += 32u - __popc(d);
                }
                if (eff >= min_identical) { myI[r]++; myJ[q]++; }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) if (myI[r]) atomicAdd(&cntI[4 * ty + r], myI[r]);
#pragma unroll
    for (int q = 0; q < 4; ++q) if (myJ[q]) atomicAdd(&cntJ[4 * tx + q], myJ[q]);
    __syncthreads();
    if (threadIdx.x < HT) {
        const uint32_t si = i0 + threadIdx.x, sj = j0 + threadIdx.x;
        if (si < n && cntI[threadIdx.x]) atomicAdd(&counts[si], (unsigned long long)cntI[threadIdx.x]);
        if (sj < n && cntJ[threadIdx.x]) atomicAdd(&counts[sj], (unsigned long long)cntJ[threadIdx.x]);
    }
}

// The reference's SAMPLE-0 double count (SPK_QUIRK_SAMPLE0): calculate_weights compares sample 0 with every sample j in
// [0, n) -- itself included -- in its cache-priming pass (range_primer, Alignment_StateVector_weights.hpp:218-238) and then
// again as pair (j, 0) of the triangular pass (range_worker, :240-262). One warp per sample j: ident(0, j) over the planes,
// the same final-block rule as above, and one more count for both ends of every similar pair (j = 0: the self pair).
__global__ void __launch_bounds__(256)
hamming_row0_kernel(const uint32_t* __restrict__ planes, uint32_t n, uint32_t L, uint32_t W32,
                    unsigned long long min_identical, int lastblock_quirk, unsigned long long* __restrict__ counts) {
    const uint32_t j = blockIdx.x * 8 + threadIdx.x / 32, lane = threadIdx.x % 32;
    if (j >= n) return;
    const size_t plane = (size_t)n * W32;
    const uint32_t* r0 = planes; const uint32_t* rj = planes + (size_t)j * W32;
    uint32_t mism = 0;
    for (uint32_t w = lane; w < W32; w += 32)
        mism += __popc((r0[w] ^ rj[w]) | (r0[plane + w] ^ rj[plane + w]) | (r0[2 * plane + w] ^ rj[2 * plane + w]));
    mism = __reduce_add_sync(0xffffffffu, mism);
    if (lane == 0) {
        unsigned long long eff = (unsigned long long)L - mism;
        if (lastblock_quirk && (L % 32u == 0u) && W32 > 0) {
            const uint32_t w = W32 - 1;
            eff += 32u - __popc((r0[w] ^ rj[w]) | (r0[plane + w] ^ rj[plane + w]) | (r0[2 * plane + w] ^ rj[2 * plane + w]));
        }
        if (eff >= min_identical) { atomicAdd(&counts[j], 1ull); atomicAdd(&counts[0], 1ull); }
    }
}

__global__ void fill_ones_kernel(unsigned long long* counts, uint32_t n, unsigned long long v) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) counts[i] = v;        // 1 = multiplicity of the sample itself (weights.hpp:205-206); partition 0 carries it
}

int hamming_neighbour_counts(Ctx* c, uint64_t min_identical, int quirks, uint32_t part, uint32_t n_parts) {
    const uint32_t nt = (c->n + HT - 1) / HT;
    const int quirk = quirks & SPK_QUIRK_LASTBLOCK;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_K1_0], c->stream));
    fill_ones_kernel<<<(c->n + 255) / 256, 256, 0, c->stream>>>((unsigned long long*)c->d_counts, c->n, part == 0 ? 1ull : 0ull);
    static const bool v1 = [] { const char* e = getenv("SPYDRPICK_B200_K1"); return e && e[0] == '0'; }();
    if (n_parts == 0) n_parts = 1;
    const unsigned long long T = (unsigned long long)nt * (nt + 1) / 2;
    if (v1 || nt >= 16384u || (unsigned long long)3 * c->n * c->W32 >= (1ull << 32))
        hamming_tile_kernel<<<dim3(nt, nt), 256, 0, c->stream>>>(c->d_planes, c->n, c->L, c->W32,
                                                                 (unsigned long long)min_identical, quirk,
                                                                 (unsigned long long*)c->d_counts, part, n_parts);
    else if (T > part) {
        // per device, so set on every launch (a process may hold contexts on several GPUs)
        SPK_CUDA(c, cudaFuncSetAttribute(hamming_csa_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, HC_SMEM_BYTES));
        const unsigned grid = (unsigned)((T - part + n_parts - 1) / n_parts);
        hamming_csa_kernel<<<grid, 256, HC_SMEM_BYTES, c->stream>>>(c->d_planes, c->n, c->L, c->W32, nt,
                                                                    (unsigned long long)min_identical, quirk,
                                                                    (unsigned long long*)c->d_counts, part, n_parts);
    }
    c->tm.kernel_launches += 2;
    if ((quirks & SPK_QUIRK_SAMPLE0) && part == 0) {           // partition 0 carries the extra counts
        hamming_row0_kernel<<<(c->n + 7) / 8, 256, 0, c->stream>>>(c->d_planes, c->n, c->L, c->W32, (unsigned long long)min_identical,
                                                                   quirk, (unsigned long long*)c->d_counts);
        c->tm.kernel_launches += 1;
    }
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_K1_1], c->stream));
    c->tm_pending |= TMP_K1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

} // namespace spk
