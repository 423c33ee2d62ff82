// spk_graph.cu -- edge sort (Graph::sort) and K3: ARACNE as a CSR sorted-intersection kernel.
//
// Graph::sort: externals/apegrunt/include/graph/Graph.hpp:234-242 -- weight desc, id asc with
// id = (node2 << 32) | node1, node1 = min, node2 = max (Graph.hpp:52-58,94-101).
// ARACNE: include/ARACNE.hpp:246-368 (block_processor::process_edge) and :411-548. The
// block-streamed reference marks, for every triangle of the final graph, each edge whose weight
// equals the triangle minimum when median - min >= threshold; marking is idempotent and never
// removes neighbours, so the result is the order-independent predicate evaluated here:
//   edge e=(u,v,w) is INDIRECT iff exists k in N(u) & N(v): mid(w,w_uk,w_vk) - min(...) >= thr and w == min.
// Each edge writes only its own flag: no atomics.
#include <cub/cub.cuh>
#include "spk_internal.h"

namespace spk {

__global__ void sort_keys_kernel(const spk_edge* __restrict__ e, unsigned long long n,
                                 unsigned long long* __restrict__ key_id, uint32_t* __restrict__ idx) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t a = min(e[k].i, e[k].j), b = max(e[k].i, e[k].j);
    key_id[k] = ((unsigned long long)b << 32) | a;
    idx[k] = (uint32_t)k;
}
__global__ void weight_keys_kernel(const spk_edge* __restrict__ e, const uint32_t* __restrict__ idx, unsigned long long n,
                                   unsigned long long* __restrict__ key_w) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    key_w[k] = ~f64_ordered(e[idx[k]].mi);           // ascending on ~ordered == descending weight
}
__global__ void gather_edges_kernel(const spk_edge* __restrict__ src, const uint32_t* __restrict__ idx,
                                    unsigned long long n, spk_edge* __restrict__ dst) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) dst[k] = src[idx[k]];
}

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

int sort_edges_device(Ctx* c, spk_edge* d_edges, uint64_t n) {
    if (n < 2) return SPK_OK;
    if (n > 0xffffffffull) return fail(c, SPK_E_INVALID, "edge sort supports < 2^32 edges (reference limit, ARACNE.hpp:416)");
    size_t tmp1 = 0, tmp2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp1, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 64, c->stream);
    tmp2 = tmp1;
    const size_t kb = align256(n * 8), ib = align256(n * 4), eb = align256(n * sizeof(spk_edge));
    uint8_t* base = (uint8_t*)scratch(c, 2 * kb + 2 * ib + eb + align256(tmp2));
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed (sort)");
    unsigned long long* k0 = (unsigned long long*)base;
    unsigned long long* k1 = (unsigned long long*)(base + kb);
    uint32_t* i0 = (uint32_t*)(base + 2 * kb);
    uint32_t* i1 = (uint32_t*)(base + 2 * kb + ib);
    spk_edge* ebuf = (spk_edge*)(base + 2 * kb + 2 * ib);
    void* tmp = base + 2 * kb + 2 * ib + eb;
    const unsigned g = (unsigned)((n + 255) / 256);
    SPK_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
    sort_keys_kernel<<<g, 256, 0, c->stream>>>(d_edges, n, k0, i0);
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmp2, k0, k1, i0, i1, (int)n, 0, 64, c->stream));   // by id asc
    weight_keys_kernel<<<g, 256, 0, c->stream>>>(d_edges, i1, n, k0);
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmp2, k0, k1, i1, i0, (int)n, 0, 64, c->stream));   // stable by weight desc
    gather_edges_kernel<<<g, 256, 0, c->stream>>>(d_edges, i0, n, ebuf);
    SPK_CUDA(c, cudaMemcpyAsync(d_edges, ebuf, n * sizeof(spk_edge), cudaMemcpyDeviceToDevice, c->stream));
    c->tm.kernel_launches += 3 + 2 * 8;     // 3 own kernels + radix passes (CUB library kernels)
    SPK_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
    SPK_CUDA(c, cudaEventSynchronize(c->ev[1]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.sort_ms, c->ev[0], c->ev[1]));
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// ---- CSR build --------------------------------------------------------------------------------
__global__ void directed_entries_kernel(const spk_edge* __restrict__ e, unsigned long long n,
                                        unsigned long long* __restrict__ keys, unsigned long long* __restrict__ vals,
                                        unsigned int* __restrict__ max_node) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int m = 0;
    if (k < n) {
        const spk_edge ed = e[k];
        keys[2 * k] = ((unsigned long long)ed.i << 32) | ed.j;
        keys[2 * k + 1] = ((unsigned long long)ed.j << 32) | ed.i;
        const unsigned long long wb = (unsigned long long)__double_as_longlong(ed.mi);
        vals[2 * k] = wb; vals[2 * k + 1] = wb;
        m = max(ed.i, ed.j);
    }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x % 32) == 0 && m) atomicMax(max_node, m);
}

// offsets[v] = first sorted entry with src >= v, offsets[nn] = 2E
__global__ void csr_offsets_kernel(const unsigned long long* __restrict__ keys, unsigned long long m,
                                   unsigned long long* __restrict__ offsets, uint32_t nn) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k > m) return;
    const long long prev = (k == 0) ? -1 : (long long)(keys[k - 1] >> 32);
    const long long cur = (k == m) ? (long long)nn : (long long)(keys[k] >> 32);
    for (long long v = prev + 1; v <= cur; ++v) offsets[v] = k;
}

// ---- K3 ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
aracne_kernel(const spk_edge* __restrict__ edges, unsigned long long e_begin, unsigned long long e_end,
              const unsigned long long* __restrict__ keys, const unsigned long long* __restrict__ vals,
              const unsigned long long* __restrict__ offsets, double threshold, uint8_t* __restrict__ flags) {
    const uint32_t lane = threadIdx.x % 32;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) / 32;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) / 32;
    for (unsigned long long e = e_begin + warp; e < e_end; e += nwarps) {
        const spk_edge ed = edges[e];
        const double w = ed.mi;
        unsigned long long a0 = offsets[ed.i], a1 = offsets[ed.i + 1];
        unsigned long long b0 = offsets[ed.j], b1 = offsets[ed.j + 1];
        if (a1 - a0 > b1 - b0) { unsigned long long t; t = a0; a0 = b0; b0 = t; t = a1; a1 = b1; b1 = t; }   // A = shorter list
        bool indirect = false;
        for (unsigned long long base = a0; base < a1 && !indirect; base += 32) {
            const unsigned long long k = base + lane;
            bool hit = false;
            if (k < a1) {
                const uint32_t node = (uint32_t)keys[k];
                // binary search node in B (sorted by dst)
                unsigned long long lo = b0, hi = b1;
                while (lo < hi) {
                    const unsigned long long mid = (lo + hi) >> 1;
                    if ((uint32_t)keys[mid] < node) lo = mid + 1; else hi = mid;
                }
                if (lo < b1 && (uint32_t)keys[lo] == node) {
                    const double w2 = __longlong_as_double((long long)vals[k]);
                    const double w3 = __longlong_as_double((long long)vals[lo]);
                    const double midval = fmax(fmin(w, w2), fmin(fmax(w, w2), w3));    // ARACNE.hpp:309
                    const double minval = fmin(w, fmin(w2, w3));                        // ARACNE.hpp:310
                    hit = (midval - minval >= threshold) && (w == minval);
                }
            }
            indirect = __any_sync(0xffffffffu, hit);
        }
        if (lane == 0) flags[e] = indirect ? 0 : 1;    // 1 = direct ("Booleans swapped", ARACNE.hpp:399-405)
    }
}

int aracne_device(Ctx* c, const spk_edge* d_edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, uint8_t* d_flags) {
    if (n == 0) return SPK_OK;
    if (n > 0x7fffffffull) return fail(c, SPK_E_INVALID, "ARACNE supports < 2^31 edges");
    const unsigned long long m = 2 * n;
    size_t tmpb = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmpb, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)m, 0, 64, c->stream);
    const size_t kb = align256(m * 8);
    // node count is unknown for caller-supplied edge lists: reserve offsets for the worst case after finding max id
    uint8_t* base = (uint8_t*)scratch(c, 4 * kb + align256(tmpb) + 256);
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed (aracne)");
    unsigned long long* k0 = (unsigned long long*)base;
    unsigned long long* k1 = (unsigned long long*)(base + kb);
    unsigned long long* v0 = (unsigned long long*)(base + 2 * kb);
    unsigned long long* v1 = (unsigned long long*)(base + 3 * kb);
    void* tmp = base + 4 * kb;
    unsigned int* d_max = (unsigned int*)(base + 4 * kb + align256(tmpb));
    SPK_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
    SPK_CUDA(c, cudaMemsetAsync(d_max, 0, 4, c->stream));
    directed_entries_kernel<<<(unsigned)((n + 255) / 256), 256, 0, c->stream>>>(d_edges, n, k0, v0, d_max);
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmpb, k0, k1, v0, v1, (int)m, 0, 64, c->stream));
    unsigned int h_max = 0;
    SPK_CUDA(c, cudaMemcpyAsync(&h_max, d_max, 4, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint32_t nn = h_max + 1;
    // offsets live in the (now free) k0 buffer: (nn+1) entries <= 2E+... guard
    unsigned long long* offsets = k0;
    unsigned long long* offsets_tmp = nullptr;
    if ((size_t)(nn + 2) * 8 > kb) {          // sparse graph on many nodes: offsets do not fit the recycled key buffer
        if (cudaMalloc(&offsets_tmp, (size_t)(nn + 2) * 8) != cudaSuccess) return fail(c, SPK_E_NOMEM, "offset table allocation failed");
        offsets = offsets_tmp;
    }
    csr_offsets_kernel<<<(unsigned)((m + 1 + 255) / 256), 256, 0, c->stream>>>(k1, m, offsets, nn);
    SPK_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
    if (nparts == 0) nparts = 1;
    const unsigned long long e0 = (unsigned long long)n * part / nparts, e1 = (unsigned long long)n * (part + 1) / nparts;
    const unsigned long long warps_needed = e1 - e0;
    unsigned long long blocks = (warps_needed + 7) / 8;
    const unsigned long long cap = (unsigned long long)c->sm_count * 32;
    if (blocks > cap) blocks = cap;
    if (blocks == 0) blocks = 1;
    aracne_kernel<<<(unsigned)blocks, 256, 0, c->stream>>>(d_edges, e0, e1, k1, v1, offsets, eps, d_flags);
    c->tm.kernel_launches += 3 + 8;
    SPK_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
    SPK_CUDA(c, cudaEventSynchronize(c->ev[2]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.aracne_build_ms, c->ev[0], c->ev[1]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.aracne_ms, c->ev[1], c->ev[2]));
    if (offsets_tmp) cudaFree(offsets_tmp);
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

} // namespace spk
