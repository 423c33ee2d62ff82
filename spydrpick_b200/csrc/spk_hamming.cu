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

__global__ void fill_ones_kernel(unsigned long long* counts, uint32_t n, unsigned long long v) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) counts[i] = v;        // 1 = multiplicity of the sample itself (weights.hpp:205-206); partition 0 carries it
}

int hamming_neighbour_counts(Ctx* c, uint64_t min_identical, int quirk, uint32_t part, uint32_t n_parts) {
    const uint32_t nt = (c->n + HT - 1) / HT;
    SPK_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
    fill_ones_kernel<<<(c->n + 255) / 256, 256, 0, c->stream>>>((unsigned long long*)c->d_counts, c->n, part == 0 ? 1ull : 0ull);
    hamming_tile_kernel<<<dim3(nt, nt), 256, 0, c->stream>>>(c->d_planes, c->n, c->L, c->W32,
                                                             (unsigned long long)min_identical, quirk,
                                                             (unsigned long long*)c->d_counts, part, n_parts);
    c->tm.kernel_launches += 2;
    SPK_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
    SPK_CUDA(c, cudaEventSynchronize(c->ev[1]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.hamming_ms, c->ev[0], c->ev[1]));
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

} // namespace spk
