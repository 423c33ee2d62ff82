// spk_noalloc.cu -- spk_neighbour_counts_noalloc: K1 (sample-pair identity counts) for callers that cannot allocate.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops, harnesses that forbid it). The regular K1 path (spk_set_alignment + spk_neighbour_counts,
// spk_prep.cu / spk_hamming.cu) keeps its bit planes in cudaMalloc'ed context memory; this variant keeps them in a file-scope
// __device__ array of SPK_NOALLOC_MAX_SCRATCH_BYTES that every GPU loading the library holds until exit. The entry point takes
// device pointers for its input (dense states) and output (counts) and a stream; it calls no cudaMalloc / cudaFree / allocating
// library routine. All calls on one device share the array: calls that overlap on the GPU race (serialise them on one stream).
//
// Same arithmetic as spk_hamming.cu: ident(i,j) = L - popc((p0_i^p0_j) | (p1_i^p1_j) | (p2_i^p2_j)) over ceil(L/32) words,
// both reference quirks (SPK_QUIRK_LASTBLOCK, SPK_QUIRK_SAMPLE0); integer work, bit-exact against spk_neighbour_counts.
#include <mutex>
#include "spk_internal.h"

namespace spk {

constexpr size_t NA_MAX_WORDS = (size_t)SPK_NOALLOC_MAX_SCRATCH_BYTES / 4;
// The scratch: [3 planes][n][W32] words for the current call. In HOST code this name is only a stand-in for the symbol, not a
// device address: the address is taken once per device with cudaGetSymbolAddress (na_scratch_ptr) and passed to the kernels.
static __device__ __align__(256) uint32_t g_na_planes[NA_MAX_WORDS];

constexpr int NA_MAX_DEVICES = 64;
static std::mutex g_na_mutex;
static uint32_t* g_na_ptr[NA_MAX_DEVICES];          // per-device address of g_na_planes

static int na_scratch_ptr(uint32_t** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(nullptr, SPK_E_CUDA, "neighbour_counts_noalloc: cudaGetDevice: %s (no usable device; there is no CPU fallback)", cudaGetErrorString(e));
    if (dev < 0 || dev >= NA_MAX_DEVICES) return fail(nullptr, SPK_E_INVALID, "neighbour_counts_noalloc: device index %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_na_mutex);
    if (!g_na_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_na_planes);         // set-up, once per device, with that device current
        if (e != cudaSuccess || !p) { cudaGetLastError(); return fail(nullptr, SPK_E_CUDA, "neighbour_counts_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); }
        g_na_ptr[dev] = (uint32_t*)p;
    }
    *out = g_na_ptr[dev];
    return SPK_OK;
}

// states[n][L] (codes 0..4) -> planes[p][s][w], bit b = bit p of state(s, 32w + b), pad bits 0. Writes every word of the
// 3 * n * W32 extent this call uses, so nothing of the previous call's contents survives inside it.
__global__ void __launch_bounds__(256)
na_pack_kernel(const uint8_t* __restrict__ states, uint32_t* __restrict__ planes, uint32_t n, uint32_t L, uint32_t W32) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)n * W32) return;
    const uint32_t s = (uint32_t)(idx / W32), w = (uint32_t)(idx % W32);
    const uint8_t* row = states + (size_t)s * L + (size_t)w * 32;
    const uint32_t cnt = min(32u, L - w * 32);
    uint32_t p0 = 0, p1 = 0, p2 = 0;
    for (uint32_t b = 0; b < cnt; ++b) {
        const uint32_t x = row[b];
        p0 |= (x & 1u) << b; p1 |= ((x >> 1) & 1u) << b; p2 |= ((x >> 2) & 1u) << b;
    }
    const size_t plane = (size_t)n * W32;
    planes[idx] = p0; planes[plane + idx] = p1; planes[2 * plane + idx] = p2;
}

__global__ void na_init_counts_kernel(unsigned long long* counts, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) counts[i] = 1ull;                      // the sample itself (Alignment_StateVector_weights.hpp:205-206)
}

constexpr int NT = 64;         // samples per tile side
constexpr int NCH = 16;        // plane words per shared-memory chunk
constexpr int NPITCH = NCH + 1;

__device__ __forceinline__ uint32_t na_lastblock_bonus(const uint32_t* __restrict__ planes, size_t plane, uint32_t W32, uint32_t a, uint32_t b) {
    const size_t oa = (size_t)a * W32 + (W32 - 1), ob = (size_t)b * W32 + (W32 - 1);
    return 32u - __popc((planes[oa] ^ planes[ob]) | (planes[plane + oa] ^ planes[plane + ob]) | (planes[2 * plane + oa] ^ planes[2 * plane + ob]));
}

// One block per upper-triangle 64 x 64 tile of sample pairs, 4 x 4 pairs per thread, planes streamed through shared memory.
__global__ void __launch_bounds__(256)
na_hamming_kernel(const uint32_t* __restrict__ planes, uint32_t n, uint32_t L, uint32_t W32, uint32_t nt,
                  unsigned long long min_identical, int lastblock, unsigned long long* __restrict__ counts) {
    // linear index over tile pairs ti <= tj: row ti starts at ti * nt - ti (ti - 1) / 2
    const uint32_t b = blockIdx.x;
    uint32_t ti = 0;
    while ((ti + 1) * nt - (ti + 1) * ti / 2 <= b) ++ti;
    const uint32_t tj = ti + (b - (ti * nt - ti * (ti - 1) / 2));
    __shared__ uint32_t sI[3][NT][NPITCH];
    __shared__ uint32_t sJ[3][NT][NPITCH];
    __shared__ unsigned int cntI[NT], cntJ[NT];
    const uint32_t tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    const uint32_t i0 = ti * NT, j0 = tj * NT;
    const size_t plane = (size_t)n * W32;
    if (threadIdx.x < NT) { cntI[threadIdx.x] = 0; cntJ[threadIdx.x] = 0; }
    uint32_t acc[4][4];
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[r][q] = 0;
    for (uint32_t w0 = 0; w0 < W32; w0 += NCH) {
        __syncthreads();
        for (uint32_t e = threadIdx.x; e < 3 * NT * NCH; e += 256) {
            const uint32_t k = e % NCH, r = (e / NCH) % NT, p = e / (NCH * NT);
            const uint32_t w = w0 + k, si = i0 + r, sj = j0 + r;
            sI[p][r][k] = (si < n && w < W32) ? planes[p * plane + (size_t)si * W32 + w] : 0u;
            sJ[p][r][k] = (sj < n && w < W32) ? planes[p * plane + (size_t)sj * W32 + w] : 0u;
        }
        __syncthreads();
#pragma unroll 4
        for (int k = 0; k < NCH; ++k) {
            uint32_t a0[4], a1[4], a2[4], b0[4], b1[4], b2[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) { a0[r] = sI[0][ty + 16 * r][k]; a1[r] = sI[1][ty + 16 * r][k]; a2[r] = sI[2][ty + 16 * r][k]; }
#pragma unroll
            for (int q = 0; q < 4; ++q) { b0[q] = sJ[0][tx + 16 * q][k]; b1[q] = sJ[1][tx + 16 * q][k]; b2[q] = sJ[2][tx + 16 * q][k]; }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[r][q] += __popc((a0[r] ^ b0[q]) | (a1[r] ^ b1[q]) | (a2[r] ^ b2[q]));
        }
    }
    const bool bonus = lastblock && (L % 32u == 0u) && W32 > 0;
    unsigned int myI[4] = {0, 0, 0, 0}, myJ[4] = {0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t si = i0 + ty + 16 * r;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const uint32_t sj = j0 + tx + 16 * q;
            if (si < n && sj < n && si < sj) {
                unsigned long long ident = (unsigned long long)L - acc[r][q];
                if (bonus) ident += na_lastblock_bonus(planes, plane, W32, si, sj);
                if (ident >= min_identical) { myI[r]++; myJ[q]++; }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) if (myI[r]) atomicAdd(&cntI[ty + 16 * r], myI[r]);
#pragma unroll
    for (int q = 0; q < 4; ++q) if (myJ[q]) atomicAdd(&cntJ[tx + 16 * q], myJ[q]);
    __syncthreads();
    if (threadIdx.x < NT) {
        const uint32_t si = i0 + threadIdx.x, sj = j0 + threadIdx.x;
        if (si < n && cntI[threadIdx.x]) atomicAdd(&counts[si], (unsigned long long)cntI[threadIdx.x]);
        if (sj < n && cntJ[threadIdx.x]) atomicAdd(&counts[sj], (unsigned long long)cntJ[threadIdx.x]);
    }
}

// SPK_QUIRK_SAMPLE0: sample 0 against every sample (itself included) once more; one warp per sample.
__global__ void __launch_bounds__(256)
na_row0_kernel(const uint32_t* __restrict__ planes, uint32_t n, uint32_t L, uint32_t W32, unsigned long long min_identical,
               int lastblock, unsigned long long* __restrict__ counts) {
    const uint32_t j = blockIdx.x * 8 + threadIdx.x / 32, lane = threadIdx.x % 32;
    if (j >= n) return;
    const size_t plane = (size_t)n * W32;
    const uint32_t* r0 = planes; const uint32_t* rj = planes + (size_t)j * W32;
    uint32_t mism = 0;
    for (uint32_t w = lane; w < W32; w += 32)
        mism += __popc((r0[w] ^ rj[w]) | (r0[plane + w] ^ rj[plane + w]) | (r0[2 * plane + w] ^ rj[2 * plane + w]));
    mism = __reduce_add_sync(0xffffffffu, mism);
    if (lane == 0) {
        unsigned long long ident = (unsigned long long)L - mism;
        if (lastblock && (L % 32u == 0u) && W32 > 0) ident += na_lastblock_bonus(planes, plane, W32, 0u, j);
        if (ident >= min_identical) { atomicAdd(&counts[j], 1ull); atomicAdd(&counts[0], 1ull); }
    }
}

} // namespace spk

using namespace spk;

extern "C" int spk_neighbour_counts_noalloc(const void* d_states, uint32_t n, uint32_t L, uint64_t min_identical, int quirks,
                                            uint64_t* d_counts, void* cuda_stream) {
    if (!d_states || !d_counts) return fail(nullptr, SPK_E_INVALID, "neighbour_counts_noalloc: NULL device pointer");
    if (n == 0 || L == 0) return fail(nullptr, SPK_E_INVALID, "neighbour_counts_noalloc: empty alignment (n = %u, L = %u)", n, L);
    const uint32_t W32 = (L + 31u) / 32u;
    const unsigned long long words = 3ull * n * W32;
    if (words > (unsigned long long)NA_MAX_WORDS)
        return fail(nullptr, SPK_E_INVALID, "neighbour_counts_noalloc: n = %u, L = %u needs %llu bytes of bit planes, more than the fixed scratch of "
                    "SPK_NOALLOC_MAX_SCRATCH_BYTES = %llu (3 * n * ceil(L / 32) * 4 must not exceed it); use spk_set_alignment + spk_neighbour_counts",
                    n, L, words * 4ull, (unsigned long long)SPK_NOALLOC_MAX_SCRATCH_BYTES);
    const unsigned long long nt = (n + NT - 1) / NT, tiles = nt * (nt + 1) / 2;
    if (tiles > 0x7fffffffull) return fail(nullptr, SPK_E_INVALID, "neighbour_counts_noalloc: too many samples");
    uint32_t* planes = nullptr;
    int r = na_scratch_ptr(&planes); if (r) return r;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int lastblock = quirks & SPK_QUIRK_LASTBLOCK;
    const size_t nw = (size_t)n * W32;
    na_pack_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, st>>>((const uint8_t*)d_states, planes, n, L, W32);
    na_init_counts_kernel<<<(n + 255) / 256, 256, 0, st>>>((unsigned long long*)d_counts, n);
    na_hamming_kernel<<<(unsigned)tiles, 256, 0, st>>>(planes, n, L, W32, (uint32_t)nt, (unsigned long long)min_identical, lastblock,
                                                        (unsigned long long*)d_counts);
    if (quirks & SPK_QUIRK_SAMPLE0)
        na_row0_kernel<<<(n + 7) / 8, 256, 0, st>>>(planes, n, L, W32, (unsigned long long)min_identical, lastblock, (unsigned long long*)d_counts);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, SPK_E_CUDA, "neighbour_counts_noalloc: launch: %s", cudaGetErrorString(e));
    return SPK_OK;
}
