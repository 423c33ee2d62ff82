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
#include <algorithm>
#include <cub/cub.cuh>
#include <vector>
#include "spk_internal.h"

namespace spk {

// id key (node2 << bits) | node1 with bits = ceil(log2(#nodes)): same order as (node2 << 32) | node1 in 2*bits radix bits
__global__ void sort_keys_kernel(const spk_edge* __restrict__ e, unsigned long long n, uint32_t bits,
                                 unsigned long long* __restrict__ key_id, uint32_t* __restrict__ idx) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t a = min(e[k].i, e[k].j), b = max(e[k].i, e[k].j);
    key_id[k] = ((unsigned long long)b << bits) | a;
    idx[k] = (uint32_t)k;
}
__global__ void max_node_kernel(const spk_edge* __restrict__ e, unsigned long long n, unsigned int* __restrict__ out) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int m = 0;
    if (k < n) m = max(e[k].i, e[k].j);
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x % 32) == 0 && m) atomicMax(out, m);
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
__global__ void widen_u32_kernel(const unsigned int* __restrict__ in, unsigned long long* __restrict__ out, uint32_t n) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = in[k];
}

int sort_edges_device(Ctx* c, spk_edge* d_edges, uint64_t n) {
    if (n < 2) return SPK_OK;
    if (n > 0x7fffffffull) return fail(c, SPK_E_INVALID, "edge sort supports < 2^31 edges (CUB item counts are int; the reference's own limit is 2^32, ARACNE.hpp:416)");
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
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_SORT0], c->stream));
    // node ids of a scan are < L: 2*ceil(log2 L) radix bits instead of 64; caller-supplied lists: largest id found first
    uint32_t nn = c->L;
    if (!c->d_states || (d_edges != c->d_edges && d_edges != c->d_edges_view)) {
        unsigned int* d_max = (unsigned int*)k1;
        SPK_CUDA(c, cudaMemsetAsync(d_max, 0, 4, c->stream));
        max_node_kernel<<<g, 256, 0, c->stream>>>(d_edges, n, d_max);
        unsigned int h_max = 0;
        SPK_CUDA(c, cudaMemcpyAsync(&h_max, d_max, 4, cudaMemcpyDeviceToHost, c->stream));
        SPK_CUDA(c, cudaStreamSynchronize(c->stream));
        nn = h_max + 1;
    }
    uint32_t bits = 1; while (bits < 32 && (1ull << bits) < (unsigned long long)nn) ++bits;
    sort_keys_kernel<<<g, 256, 0, c->stream>>>(d_edges, n, bits, k0, i0);
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmp2, k0, k1, i0, i1, (int)n, 0, (int)(2 * bits), c->stream));   // by id asc
    weight_keys_kernel<<<g, 256, 0, c->stream>>>(d_edges, i1, n, k0);
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmp2, k0, k1, i1, i0, (int)n, 0, 64, c->stream));   // stable by weight desc
    gather_edges_kernel<<<g, 256, 0, c->stream>>>(d_edges, i0, n, ebuf);
    SPK_CUDA(c, cudaMemcpyAsync(d_edges, ebuf, n * sizeof(spk_edge), cudaMemcpyDeviceToDevice, c->stream));
    c->tm.kernel_launches += 3 + 2 * 8;     // 3 own kernels + radix passes (CUB library kernels)
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_SORT1], c->stream));
    c->tm_pending |= TMP_SORT;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// ---- merge of sorted runs (multi-GPU: every rank sorts its own edges, the all-gathered list is a concatenation of G
// sorted runs) -- Graph::sort's order (weight descending, id ascending; graph/Graph.hpp:234-242) is a strict total
// order on distinct edges, so a pairwise merge tree gives exactly the list a full sort gives.
struct EdgeBefore {
    __host__ __device__ bool operator()(const spk_edge& a, const spk_edge& b) const {
        if (a.mi != b.mi) return a.mi > b.mi;
        const unsigned long long ia = ((unsigned long long)a.j << 32) | a.i, ib = ((unsigned long long)b.j << 32) | b.i;
        return ia < ib;
    }
};

int merge_sorted_runs_device(Ctx* c, spk_edge* d_edges, const uint64_t* run_counts, uint32_t n_runs) {
    std::vector<uint64_t> off(1, 0);
    for (uint32_t r = 0; r < n_runs; ++r) if (run_counts[r]) off.push_back(off.back() + run_counts[r]);
    const uint64_t n = off.back();
    if (off.size() <= 2) return SPK_OK;                               // zero or one non-empty run
    if (n > 0x7fffffffull) return fail(c, SPK_E_INVALID, "run merge supports < 2^31 edges");
    size_t tmpb = 0;
    cub::DeviceMerge::MergeKeys(nullptr, tmpb, (const spk_edge*)nullptr, (int)n, (const spk_edge*)nullptr, (int)n, (spk_edge*)nullptr, EdgeBefore{}, c->stream);
    const size_t eb = align256(n * sizeof(spk_edge));
    uint8_t* base = (uint8_t*)scratch(c, eb + align256(tmpb) + 256);
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed (run merge)");
    spk_edge* src = d_edges; spk_edge* dst = (spk_edge*)base;
    void* tmp = base + eb;
    while (off.size() > 2) {
        std::vector<uint64_t> next(1, 0);
        for (size_t r = 0; r + 1 < off.size(); r += 2) {
            const uint64_t a0 = off[r], a1 = off[r + 1];
            if (r + 2 < off.size()) {
                const uint64_t b1 = off[r + 2];
                size_t tb = tmpb;
                SPK_CUDA(c, cub::DeviceMerge::MergeKeys(tmp, tb, (const spk_edge*)(src + a0), (int)(a1 - a0), (const spk_edge*)(src + a1), (int)(b1 - a1),
                                                        dst + a0, EdgeBefore{}, c->stream));
                c->tm.kernel_launches += 2;
                next.push_back(b1);
            } else {                                                   // odd run out: carried to the next round
                SPK_CUDA(c, cudaMemcpyAsync(dst + a0, src + a0, (a1 - a0) * sizeof(spk_edge), cudaMemcpyDeviceToDevice, c->stream));
                next.push_back(a1);
            }
        }
        off.swap(next);
        std::swap(src, dst);
    }
    if (src != d_edges) SPK_CUDA(c, cudaMemcpyAsync(d_edges, src, n * sizeof(spk_edge), cudaMemcpyDeviceToDevice, c->stream));
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

// ---- adjacency build --------------------------------------------------------------------------------
// Over the 2E directed entries (entry 2r / 2r+1 = edge r seen from i / from j):
//   offsets[]            per node, from a degree histogram + exclusive scan;
//   wdst[], wval[], weid[]  per node its neighbours in WEIGHT-DESCENDING order (the input is Graph::sort order, so ONE stable
//                        radix sort of the entries by source node -- ceil(log2 nodes) key bits -- keeps every list ordered);
//   hkey[], hval[]       per node an open-addressing hash table neighbour -> weight in the slot range [2*offsets[u], 2*offsets[u+1])
//                        (load factor 1/2), which replaces the id-ordered lists + binary searches of the first version (a second,
//                        2*ceil(log2 nodes)-bit sort and deg*log(deg) divergent loads per edge).
__global__ void directed_src_kernel(const spk_edge* __restrict__ e, unsigned long long n, uint32_t* __restrict__ wkeys, uint32_t* __restrict__ ents,
                                    unsigned int* __restrict__ degree) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const spk_edge ed = e[k];
    wkeys[2 * k] = ed.i; wkeys[2 * k + 1] = ed.j;
    ents[2 * k] = (uint32_t)(2 * k); ents[2 * k + 1] = (uint32_t)(2 * k + 1);
    atomicAdd(&degree[ed.i], 1u); atomicAdd(&degree[ed.j], 1u);
}
// out[0] = max node id, out[2] = 1 if the list is not sorted by weight descending
__global__ void edge_scan_kernel(const spk_edge* __restrict__ e, unsigned long long n, unsigned int* __restrict__ out) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int m = 0; bool bad = false;
    if (k < n) {
        m = max(e[k].i, e[k].j);
        if (k + 1 < n) bad = e[k].mi < e[k + 1].mi;
    }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    bad = __any_sync(0xffffffffu, bad);
    if ((threadIdx.x % 32) == 0) { if (m) atomicMax(&out[0], m); if (bad) out[2] = 1u; }
}
// device words -> mapped host memory (visible to the host once the stream has drained)
__global__ void publish_words_kernel(const unsigned int* __restrict__ src, unsigned int* __restrict__ dst, int nwords) {
    if ((int)threadIdx.x < nwords) dst[threadIdx.x] = src[threadIdx.x];
    __threadfence_system();
}
// entry order for an unsorted list: key = ~ordered(weight) so that an ascending stable sort puts strong edges first
__global__ void entry_weight_keys_kernel(const spk_edge* __restrict__ e, unsigned long long m,
                                         unsigned long long* __restrict__ keys, uint32_t* __restrict__ ents) {
    const unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= m) return;
    keys[q] = ~f64_ordered(e[q >> 1].mi); ents[q] = (uint32_t)q;
}
__global__ void entry_src_keys_kernel(const spk_edge* __restrict__ e, const uint32_t* __restrict__ ents, unsigned long long m,
                                      uint32_t* __restrict__ wkeys) {
    const unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= m) return;
    const uint32_t en = ents[q];
    wkeys[q] = (en & 1u) ? e[en >> 1].j : e[en >> 1].i;
}
__device__ __forceinline__ uint32_t nbr_hash(uint32_t z) { z *= 0x9E3779B1u; z ^= z >> 15; return z; }
// weight-descending lists + the per-node hash tables, from the entries sorted by source node
__global__ void lists_and_hash_kernel(const spk_edge* __restrict__ e, const uint32_t* __restrict__ ents, const uint32_t* __restrict__ src_sorted,
                                      unsigned long long m, const unsigned long long* __restrict__ offsets,
                                      uint32_t* __restrict__ wdst, double* __restrict__ wval,
                                      uint32_t* __restrict__ hkey, double* __restrict__ hval) {
    const unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= m) return;
    const uint32_t en = ents[q];
    const spk_edge ed = e[en >> 1];
    const uint32_t dst = (en & 1u) ? ed.i : ed.j, src = src_sorted[q];
    wdst[q] = dst; wval[q] = ed.mi;
    const unsigned long long t0 = 2 * offsets[src];
    const uint32_t cap = (uint32_t)(2 * (offsets[src + 1] - offsets[src]));
    uint32_t slot = (uint32_t)(((unsigned long long)nbr_hash(dst) * cap) >> 32);
    while (atomicCAS(&hkey[t0 + slot], 0xffffffffu, dst) != 0xffffffffu) { if (++slot == cap) slot = 0; }
    hval[t0 + slot] = ed.mi;
}

// ---- K3 ------------------------------------------------------------------------------------------
__device__ __forceinline__ bool aracne_hit(double w, double w2, double w3, double threshold) {
    const double midval = fmax(fmin(w, w2), fmin(fmax(w, w2), w3));    // ARACNE.hpp:309
    const double minval = fmin(w, fmin(w2, w3));                        // ARACNE.hpp:310
    return (midval - minval >= threshold) && (w == minval);
}
// weight of edge (u, z) from u's hash table [t0, t0 + cap); z must be a neighbour of u (the caller checked the bitmap)
__device__ __forceinline__ double nbr_weight(const uint32_t* __restrict__ hkey, const double* __restrict__ hval, unsigned long long t0, uint32_t cap, uint32_t z) {
    uint32_t slot = (uint32_t)(((unsigned long long)nbr_hash(z) * cap) >> 32);
    for (uint32_t probes = 0; probes < cap; ++probes) {
        const uint32_t k = hkey[t0 + slot];
        if (k == z) return hval[t0 + slot];
        if (k == 0xffffffffu) break;
        if (++slot == cap) slot = 0;
    }
    return __longlong_as_double(0x7ff8000000000000ll);               // not a neighbour (cannot happen after the bitmap test)
}

// One block per node u (dynamic queue). u's neighbourhood is a bitmap in shared memory; every edge (u,v) OWNED by u (the
// endpoint with the larger degree) is given to a warp that streams v's weight-descending list (coalesced), stops at the first
// weight < w (no triangle with w as its minimum beyond that) and looks the members that are also neighbours of u up in u's
// hash table (hot in L1: a few KB per node).
__global__ void __launch_bounds__(256)
aracne_hash_kernel(unsigned long long e_begin, unsigned long long e_end, uint32_t node_part, uint32_t node_nparts,
                   const uint32_t* __restrict__ wdst, const double* __restrict__ wval, const uint32_t* __restrict__ weid,
                   const uint32_t* __restrict__ hkey, const double* __restrict__ hval,
                   const unsigned long long* __restrict__ offsets, uint32_t nn, unsigned int* __restrict__ next_node,
                   double threshold, uint8_t* __restrict__ flags) {
    extern __shared__ uint32_t bm[];
    __shared__ uint32_t s_u;
    const uint32_t lane = threadIdx.x % 32, wid = threadIdx.x / 32, nw = blockDim.x / 32;
    for (uint32_t k = threadIdx.x; k < (nn + 31) / 32; k += blockDim.x) bm[k] = 0u;
    while (true) {
        __syncthreads();
        if (threadIdx.x == 0) s_u = atomicAdd(next_node, 1u);
        __syncthreads();
        // node shard (collective form): this partition visits nodes node_part, node_part + node_nparts, ...; every edge is owned by
        // exactly one of its endpoints, so the node shards partition the edge set and each rank pays the bitmap work of its nodes only
        const unsigned long long u64 = (unsigned long long)s_u * node_nparts + node_part;
        if (u64 >= nn) break;
        const uint32_t u = (uint32_t)u64;
        const unsigned long long a0 = offsets[u], a1 = offsets[u + 1], degu = a1 - a0;
        if (degu == 0) continue;
        for (unsigned long long k = a0 + threadIdx.x; k < a1; k += blockDim.x) { const uint32_t z = wdst[k]; atomicOr(&bm[z >> 5], 1u << (z & 31u)); }
        __syncthreads();
        const unsigned long long t0 = 2 * a0; const uint32_t cap = (uint32_t)(2 * degu);
        for (unsigned long long k = a0 + wid; k < a1; k += nw) {
            const uint32_t v = wdst[k];
            const unsigned long long e = weid[k] >> 1;
            if (e < e_begin || e >= e_end) continue;
            const unsigned long long b0 = offsets[v], b1 = offsets[v + 1], degv = b1 - b0;
            if (degv > degu || (degv == degu && v > u)) continue;             // the other endpoint owns this edge
            const double w = wval[k];
            bool indirect = false, done = false;
            for (unsigned long long base = b0; base < b1 && !done; base += 32) {
                const unsigned long long q = base + lane;
                bool hit = false, below = false;
                if (q < b1) {
                    const double w2 = wval[q];
                    below = w2 < w;
                    if (!below) {
                        const uint32_t z = wdst[q];
                        if ((bm[z >> 5] >> (z & 31u)) & 1u) hit = aracne_hit(w, w2, nbr_weight(hkey, hval, t0, cap, z), threshold);
                    }
                }
                indirect = __any_sync(0xffffffffu, hit);
                done = indirect || __any_sync(0xffffffffu, below);
            }
            if (lane == 0) flags[e] = indirect ? 0 : 1;    // 1 = direct ("Booleans swapped", ARACNE.hpp:399-405)
        }
        __syncthreads();
        for (unsigned long long k = a0 + threadIdx.x; k < a1; k += blockDim.x) { const uint32_t z = wdst[k]; bm[z >> 5] = 0u; }
    }
}

// Fallback for graphs on more nodes than a shared-memory bitmap holds: warp per edge, walk the list of the lower-degree
// endpoint, look every member with weight >= w up in the other endpoint's hash table.
__global__ void __launch_bounds__(256)
aracne_search_kernel(const spk_edge* __restrict__ edges, unsigned long long e_begin, unsigned long long e_end,
                     const uint32_t* __restrict__ wdst, const double* __restrict__ wval,
                     const uint32_t* __restrict__ hkey, const double* __restrict__ hval,
                     const unsigned long long* __restrict__ offsets, double threshold, uint8_t* __restrict__ flags) {
    const uint32_t lane = threadIdx.x % 32;
    const unsigned long long warp = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) / 32;
    const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) / 32;
    for (unsigned long long e = e_begin + warp; e < e_end; e += nwarps) {
        const spk_edge ed = edges[e];
        const double w = ed.mi;
        uint32_t u = ed.i, v = ed.j;
        if (offsets[u + 1] - offsets[u] < offsets[v + 1] - offsets[v]) { const uint32_t t = u; u = v; v = t; }      // v = shorter list
        const unsigned long long b0 = offsets[v], b1 = offsets[v + 1];
        const unsigned long long t0 = 2 * offsets[u]; const uint32_t cap = (uint32_t)(2 * (offsets[u + 1] - offsets[u]));
        bool indirect = false, done = false;
        for (unsigned long long base = b0; base < b1 && !done; base += 32) {
            const unsigned long long q = base + lane;
            bool hit = false, below = false;
            if (q < b1) {
                const double w2 = wval[q];
                below = w2 < w;
                if (!below) { const double w3 = nbr_weight(hkey, hval, t0, cap, wdst[q]); if (w3 == w3) hit = aracne_hit(w, w2, w3, threshold); }
            }
            indirect = __any_sync(0xffffffffu, hit);
            done = indirect || __any_sync(0xffffffffu, below);
        }
        if (lane == 0) flags[e] = indirect ? 0 : 1;
    }
}

// by_node = false: the partition is the edge range [part*n/nparts, (part+1)*n/nparts) (the public part_index / n_parts contract);
// by_node = true (collective form): the partition is the set of edges owned by nodes u = part (mod nparts) -- some disjoint
// cover of the edge list, combined by the caller with an element-wise max over zero-initialised flags.
int aracne_device(Ctx* c, const spk_edge* d_edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, uint8_t* d_flags, bool by_node) {
    if (n == 0) return SPK_OK;
    if (n > 0x3fffffffull) return fail(c, SPK_E_INVALID, "ARACNE supports < 2^30 edges");
    const unsigned long long m = 2 * n;
    size_t tmpb = 0, tmpb2 = 0, tmpb3 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmpb, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (uint32_t*)nullptr, (uint32_t*)nullptr, (int)m, 0, 64, c->stream);
    cub::DeviceRadixSort::SortPairs(nullptr, tmpb2, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (uint32_t*)nullptr, (int)m, 0, 32, c->stream);
    if (tmpb2 > tmpb) tmpb = tmpb2;
    const unsigned ge = (unsigned)((n + 255) / 256), gm = (unsigned)((m + 255) / 256);
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AR0], c->stream));
    // node count is unknown for caller-supplied edge lists: found on the device first (one small sync). The last scan's own
    // sorted list needs neither (node ids < L, order known) -- and must not: the 16-byte read-back would queue on the D2H
    // copy engine behind the asynchronous download of that very list (measured: CSR build 3.0 -> 5.7 ms).
    unsigned int h_scan[4] = {0, 0, 0, 0};
    const bool own = (d_edges == c->d_edges || d_edges == c->d_edges_view) && c->d_edges_sorted_n == n && c->d_states;
    if (own) {
        h_scan[0] = c->L - 1;
    } else {
        // caller-supplied list: the 16 bytes come back through a mapped page-locked mailbox written by a kernel, not through
        // cudaMemcpy -- a caller that downloads the list on another stream meanwhile (multi-GPU rank 0) owns the copy engine
        if (!c->h_mail) {
            if (cudaHostAlloc((void**)&c->h_mail, 64, cudaHostAllocMapped) != cudaSuccess ||
                cudaHostGetDevicePointer((void**)&c->d_mail, c->h_mail, 0) != cudaSuccess) {
                cudaGetLastError(); if (c->h_mail) { cudaFreeHost(c->h_mail); c->h_mail = nullptr; }
                return fail(c, SPK_E_NOMEM, "mapped mailbox allocation failed");
            }
        }
        unsigned int* d_pre = (unsigned int*)scratch(c, 256);
        if (!d_pre) return fail(c, SPK_E_NOMEM, "scratch allocation failed (aracne)");
        SPK_CUDA(c, cudaMemsetAsync(d_pre, 0, 16, c->stream));
        edge_scan_kernel<<<ge, 256, 0, c->stream>>>(d_edges, n, d_pre);
        publish_words_kernel<<<1, 32, 0, c->stream>>>(d_pre, c->d_mail, 4);
        SPK_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int k = 0; k < 4; ++k) h_scan[k] = ((volatile unsigned int*)c->h_mail)[k];
    }
    const uint32_t nn = h_scan[0] + 1;
    uint32_t bits = 1; while (bits < 32 && (1ull << bits) < (unsigned long long)nn) ++bits;
    cub::DeviceScan::ExclusiveSum(nullptr, tmpb3, (unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)(nn + 1), c->stream);
    if (tmpb3 > tmpb) tmpb = tmpb3;
    const size_t kb = align256(m * 8), hb = align256(m * 4), ob = align256((size_t)(nn + 2) * 8);
    // layout: [wval kb][hval 2kb][k64 tmp keys 2kb, only for unsorted input][wk0 wk1 we0 we1 wdst : 5hb][hkey 2hb][deg64 ob][offsets ob][degree ob][tmp][scan words]
    const size_t k64b = h_scan[2] ? 2 * kb : 0;
    uint8_t* base = (uint8_t*)scratch(c, 3 * kb + k64b + 7 * hb + 3 * ob + align256(tmpb) + 256);
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed (aracne)");
    double* wval = (double*)base;
    double* hval = (double*)(base + kb);
    unsigned long long* k64a = (unsigned long long*)(base + 3 * kb);
    unsigned long long* k64b_ = (unsigned long long*)(base + 3 * kb + kb);
    uint8_t* hbase = base + 3 * kb + k64b;
    uint32_t* wk0 = (uint32_t*)(hbase);
    uint32_t* wk1 = (uint32_t*)(hbase + hb);
    uint32_t* we0 = (uint32_t*)(hbase + 2 * hb);
    uint32_t* we1 = (uint32_t*)(hbase + 3 * hb);
    uint32_t* wdst = (uint32_t*)(hbase + 4 * hb);
    uint32_t* hkey = (uint32_t*)(hbase + 5 * hb);
    unsigned long long* deg64 = (unsigned long long*)(hbase + 7 * hb);
    unsigned long long* offsets = (unsigned long long*)(hbase + 7 * hb + ob);
    unsigned int* degree = (unsigned int*)(hbase + 7 * hb + 2 * ob);     // u32 histogram (its own region: a sparse graph has more nodes than entries)
    void* tmp = hbase + 7 * hb + 3 * ob;
    unsigned int* d_scan = (unsigned int*)(hbase + 7 * hb + 3 * ob + align256(tmpb));   // [1] node queue
    SPK_CUDA(c, cudaMemsetAsync(d_scan, 0, 16, c->stream));
    SPK_CUDA(c, cudaMemsetAsync(degree, 0, (size_t)(nn + 1) * 4, c->stream));
    SPK_CUDA(c, cudaMemsetAsync(hkey, 0xff, (size_t)m * 2 * 4, c->stream));
    directed_src_kernel<<<ge, 256, 0, c->stream>>>(d_edges, n, wk0, we0, degree);
    // degrees -> u64 -> exclusive scan (offsets[nn] = 2E)
    widen_u32_kernel<<<(nn + 1 + 255) / 256, 256, 0, c->stream>>>(degree, deg64, nn + 1);
    SPK_CUDA(c, cub::DeviceScan::ExclusiveSum(tmp, tmpb, deg64, offsets, (int)(nn + 1), c->stream));
    // weight-descending lists: the input is sorted by weight (Graph::sort), so a STABLE sort of the entries by source node
    // alone keeps them in weight order inside every list; an unsorted caller-supplied list is ordered by weight first
    uint32_t* went_in = we0;
    if (h_scan[2]) {
        entry_weight_keys_kernel<<<gm, 256, 0, c->stream>>>(d_edges, m, k64a, we1);
        SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmpb, k64a, k64b_, we1, we0, (int)m, 0, 64, c->stream));
        entry_src_keys_kernel<<<gm, 256, 0, c->stream>>>(d_edges, we0, m, wk0);
        c->tm.kernel_launches += 2 + 8;
    }
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmpb, wk0, wk1, went_in, we1, (int)m, 0, (int)bits, c->stream));    // stable by src
    lists_and_hash_kernel<<<gm, 256, 0, c->stream>>>(d_edges, we1, wk1, m, offsets, wdst, wval, hkey, hval);
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AR1], c->stream));
    if (nparts == 0) nparts = 1;
    unsigned long long e0 = (unsigned long long)n * part / nparts, e1 = (unsigned long long)n * (part + 1) / nparts;
    const size_t bm_bytes = (size_t)((nn + 31) / 32) * 4;
    const bool node_shard = by_node && nparts > 1 && bm_bytes <= 160 * 1024;
    if (node_shard) { e0 = 0; e1 = n; }
    if (bm_bytes <= 160 * 1024) {
        if (bm_bytes > 48 * 1024) SPK_CUDA(c, cudaFuncSetAttribute(aracne_hash_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bm_bytes));
        const unsigned per_sm = (unsigned)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / (bm_bytes + 1024)));
        unsigned long long blocks = (unsigned long long)c->sm_count * per_sm;
        if (blocks > nn) blocks = nn;
        aracne_hash_kernel<<<(unsigned)blocks, 256, bm_bytes, c->stream>>>(e0, e1, node_shard ? part : 0u, node_shard ? nparts : 1u, wdst, wval, we1, hkey, hval, offsets, nn, d_scan + 1, eps, d_flags);
    } else {
        unsigned long long blocks = (e1 - e0 + 7) / 8;
        const unsigned long long cap = (unsigned long long)c->sm_count * 32;
        if (blocks > cap) blocks = cap;
        if (blocks == 0) blocks = 1;
        aracne_search_kernel<<<(unsigned)blocks, 256, 0, c->stream>>>(d_edges, e0, e1, wdst, wval, hkey, hval, offsets, eps, d_flags);
    }
    c->tm.kernel_launches += 6 + (bits + 7) / 8;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AR2], c->stream));
    c->tm_pending |= TMP_ARB | TMP_AR;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

} // namespace spk
