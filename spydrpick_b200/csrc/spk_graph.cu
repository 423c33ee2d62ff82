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

// ---- CSR build --------------------------------------------------------------------------------
// One adjacency structure over the 2E directed entries (entry 2r / 2r+1 = edge r seen from i / from j), per node
// sorted by neighbour id: idst[] neighbour, ival[] weight, ieid[] edge index; offsets[] per node.
__global__ void directed_entries_kernel(const spk_edge* __restrict__ e, unsigned long long n, uint32_t bits,
                                        unsigned long long* __restrict__ keys, uint32_t* __restrict__ ents,
                                        uint32_t* __restrict__ wkeys) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const spk_edge ed = e[k];
    keys[2 * k] = ((unsigned long long)ed.i << bits) | ed.j;
    keys[2 * k + 1] = ((unsigned long long)ed.j << bits) | ed.i;
    ents[2 * k] = (uint32_t)(2 * k); ents[2 * k + 1] = (uint32_t)(2 * k + 1);
    wkeys[2 * k] = ed.i; wkeys[2 * k + 1] = ed.j;
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
// weight-descending lists (same offsets as the id-ordered ones): neighbour and weight per entry
__global__ void weight_lists_kernel(const spk_edge* __restrict__ e, const uint32_t* __restrict__ ents, unsigned long long m,
                                    uint32_t* __restrict__ wdst, double* __restrict__ wval) {
    const unsigned long long q = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= m) return;
    const uint32_t en = ents[q];
    const spk_edge ed = e[en >> 1];
    wdst[q] = (en & 1u) ? ed.i : ed.j;
    wval[q] = ed.mi;
}
// offsets[v] = first sorted entry with src >= v, offsets[nn] = 2E; neighbour / weight / edge index per sorted entry
__global__ void csr_finish_kernel(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ ents,
                                  const spk_edge* __restrict__ e, unsigned long long m, uint32_t bits,
                                  unsigned long long* __restrict__ offsets, uint32_t nn,
                                  uint32_t* __restrict__ idst, double* __restrict__ ival, uint32_t* __restrict__ ieid) {
    const unsigned long long k = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k > m) return;
    const long long prev = (k == 0) ? -1 : (long long)(keys[k - 1] >> bits);
    const long long cur = (k == m) ? (long long)nn : (long long)(keys[k] >> bits);
    for (long long v = prev + 1; v <= cur; ++v) offsets[v] = k;
    if (k < m) {
        const uint32_t en = ents[k];
        idst[k] = (uint32_t)(keys[k] & ((1ull << bits) - 1ull));
        ival[k] = e[en >> 1].mi;
        ieid[k] = en >> 1;
    }
}

// ---- K3 ------------------------------------------------------------------------------------------
__device__ __forceinline__ bool aracne_hit(double w, double w2, double w3, double threshold) {
    const double midval = fmax(fmin(w, w2), fmin(fmax(w, w2), w3));    // ARACNE.hpp:309
    const double minval = fmin(w, fmin(w2, w3));                        // ARACNE.hpp:310
    return (midval - minval >= threshold) && (w == minval);
}
// position of `node` in the id-ordered list [b0, b1), or b1
__device__ __forceinline__ unsigned long long find_neighbour(const uint32_t* __restrict__ idst, unsigned long long b0, unsigned long long b1, uint32_t node) {
    unsigned long long lo = b0, hi = b1;
    while (lo < hi) { const unsigned long long mid = (lo + hi) >> 1; if (idst[mid] < node) lo = mid + 1; else hi = mid; }
    return (lo < b1 && idst[lo] == node) ? lo : b1;
}

// Main kernel: one block per node u (dynamic queue). u's neighbourhood is a bitmap in shared memory; every edge (u,v)
// OWNED by u (u = the endpoint with the larger degree) is given to a warp that streams v's list (coalesced) and looks
// only the members that are also neighbours of u, with both triangle sides >= w, up in u's list. The generic
// formulation (walk the shorter list, binary-search every member in the longer one: aracne_search_kernel below) does
// deg x log(deg) divergent loads per edge.
__global__ void __launch_bounds__(256)
aracne_bitmap_kernel(unsigned long long e_begin, unsigned long long e_end,
                     const uint32_t* __restrict__ idst, const double* __restrict__ ival, const uint32_t* __restrict__ ieid,
                     const uint32_t* __restrict__ wdst, const double* __restrict__ wval,
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
        const uint32_t u = s_u;
        if (u >= nn) break;
        const unsigned long long a0 = offsets[u], a1 = offsets[u + 1], degu = a1 - a0;
        if (degu == 0) continue;
        for (unsigned long long k = a0 + threadIdx.x; k < a1; k += blockDim.x) { const uint32_t z = idst[k]; atomicOr(&bm[z >> 5], 1u << (z & 31u)); }
        __syncthreads();
        for (unsigned long long k = a0 + wid; k < a1; k += nw) {
            const uint32_t v = idst[k];
            const unsigned long long e = ieid[k];
            if (e < e_begin || e >= e_end) continue;
            const unsigned long long b0 = offsets[v], b1 = offsets[v + 1], degv = b1 - b0;
            if (degv > degu || (degv == degu && v > u)) continue;             // the other endpoint owns this edge
            const double w = ival[k];
            // v's neighbours by weight descending: only the prefix with w_vz >= w can close a triangle in which w is the
            // minimum, and strong sides (the likely witnesses) come first
            bool indirect = false, done = false;
            for (unsigned long long base = b0; base < b1 && !done; base += 32) {
                const unsigned long long q = base + lane;
                bool hit = false, below = false;
                if (q < b1) {
                    const double w2 = wval[q];
                    below = w2 < w;
                    if (!below) {
                        const uint32_t z = wdst[q];
                        if ((bm[z >> 5] >> (z & 31u)) & 1u) {
                            const unsigned long long pz = find_neighbour(idst, a0, a1, z);
                            if (pz < a1) hit = aracne_hit(w, w2, ival[pz], threshold);
                        }
                    }
                }
                indirect = __any_sync(0xffffffffu, hit);
                done = indirect || __any_sync(0xffffffffu, below);
            }
            if (lane == 0) flags[e] = indirect ? 0 : 1;    // 1 = direct ("Booleans swapped", ARACNE.hpp:399-405)
        }
        __syncthreads();
        for (unsigned long long k = a0 + threadIdx.x; k < a1; k += blockDim.x) { const uint32_t z = idst[k]; bm[z >> 5] = 0u; }
    }
}

// Fallback for graphs on more nodes than a shared-memory bitmap holds: warp per edge, walk the shorter list.
__global__ void __launch_bounds__(256)
aracne_search_kernel(const spk_edge* __restrict__ edges, unsigned long long e_begin, unsigned long long e_end,
                     const uint32_t* __restrict__ idst, const double* __restrict__ ival,
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
                const double w2 = ival[k];
                if (w2 >= w) {
                    const unsigned long long pz = find_neighbour(idst, b0, b1, idst[k]);
                    if (pz < b1) hit = aracne_hit(w, w2, ival[pz], threshold);
                }
            }
            indirect = __any_sync(0xffffffffu, hit);
        }
        if (lane == 0) flags[e] = indirect ? 0 : 1;
    }
}

int aracne_device(Ctx* c, const spk_edge* d_edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, uint8_t* d_flags) {
    if (n == 0) return SPK_OK;
    if (n > 0x7fffffffull) return fail(c, SPK_E_INVALID, "ARACNE supports < 2^31 edges");
    const unsigned long long m = 2 * n;
    size_t tmpb = 0, tmpb2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmpb, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                    (uint32_t*)nullptr, (uint32_t*)nullptr, (int)m, 0, 64, c->stream);
    cub::DeviceRadixSort::SortPairs(nullptr, tmpb2, (uint32_t*)nullptr, (uint32_t*)nullptr,
                                    (uint32_t*)nullptr, (uint32_t*)nullptr, (int)m, 0, 32, c->stream);
    if (tmpb2 > tmpb) tmpb = tmpb2;
    const size_t kb = align256(m * 8), hb = align256(m * 4);
    uint8_t* base = (uint8_t*)scratch(c, 4 * kb + 9 * hb + align256(tmpb) + 256);
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed (aracne)");
    unsigned long long* k0 = (unsigned long long*)base;                 // sort keys in, later offsets
    unsigned long long* k1 = (unsigned long long*)(base + kb);          // sort keys out
    double* ival = (double*)(base + 2 * kb);
    double* wval = (double*)(base + 3 * kb);
    uint8_t* hbase = base + 4 * kb;
    uint32_t* e0buf = (uint32_t*)(hbase);
    uint32_t* e1buf = (uint32_t*)(hbase + hb);
    uint32_t* idst = (uint32_t*)(hbase + 2 * hb);
    uint32_t* ieid = (uint32_t*)(hbase + 3 * hb);
    uint32_t* wk0 = (uint32_t*)(hbase + 4 * hb);
    uint32_t* wk1 = (uint32_t*)(hbase + 5 * hb);
    uint32_t* we0 = (uint32_t*)(hbase + 6 * hb);
    uint32_t* we1 = (uint32_t*)(hbase + 7 * hb);
    uint32_t* wdst = (uint32_t*)(hbase + 8 * hb);
    void* tmp = hbase + 9 * hb;
    unsigned int* d_scan = (unsigned int*)(hbase + 9 * hb + align256(tmpb));   // [0] max node id, [1] node queue, [2] unsorted flag
    const unsigned ge = (unsigned)((n + 255) / 256), gm = (unsigned)((m + 255) / 256);
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AR0], c->stream));
    // node count is unknown for caller-supplied edge lists: found on the device first (one small sync). The last scan's own
    // sorted list needs neither (node ids < L, order known) -- and must not: the 16-byte read-back would queue on the D2H
    // copy engine behind the asynchronous download of that very list (measured: CSR build 3.0 -> 5.7 ms).
    SPK_CUDA(c, cudaMemsetAsync(d_scan, 0, 16, c->stream));
    unsigned int h_scan[4] = {0, 0, 0, 0};
    if ((d_edges == c->d_edges || d_edges == c->d_edges_view) && c->d_edges_sorted_n == n && c->d_states) {
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
        edge_scan_kernel<<<ge, 256, 0, c->stream>>>(d_edges, n, d_scan);
        publish_words_kernel<<<1, 32, 0, c->stream>>>(d_scan, c->d_mail, 4);
        SPK_CUDA(c, cudaStreamSynchronize(c->stream));
        for (int k = 0; k < 4; ++k) h_scan[k] = ((volatile unsigned int*)c->h_mail)[k];
    }
    const uint32_t nn = h_scan[0] + 1;
    uint32_t bits = 1; while (bits < 32 && (1ull << bits) < (unsigned long long)nn) ++bits;
    directed_entries_kernel<<<ge, 256, 0, c->stream>>>(d_edges, n, bits, k0, e0buf, wk0);
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmpb, k0, k1, e0buf, e1buf, (int)m, 0, (int)(2 * bits), c->stream));   // (src, dst) ascending
    // weight-descending lists: the input is sorted by weight (Graph::sort), so a STABLE sort of the entries by source node
    // alone keeps them in weight order inside every list; an unsorted caller-supplied list is ordered by weight first
    uint32_t* went_in = e0buf;                                           // identity entry order (still intact: sort input)
    if (h_scan[2]) {
        entry_weight_keys_kernel<<<gm, 256, 0, c->stream>>>(d_edges, m, k0, we1);
        SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmpb, k0, (unsigned long long*)wval, we1, we0, (int)m, 0, 64, c->stream));
        entry_src_keys_kernel<<<gm, 256, 0, c->stream>>>(d_edges, we0, m, wk0);
        went_in = we0;
        c->tm.kernel_launches += 2 + 8;
    }
    SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(tmp, tmpb, wk0, wk1, went_in, we1, (int)m, 0, (int)bits, c->stream));    // stable by src
    // offsets live in the (now free) k0 buffer when they fit
    unsigned long long* offsets = k0;
    unsigned long long* offsets_tmp = nullptr;
    if ((size_t)(nn + 2) * 8 > kb) {          // sparse graph on many nodes: offsets do not fit the recycled key buffer
        if (cudaMalloc(&offsets_tmp, (size_t)(nn + 2) * 8) != cudaSuccess) return fail(c, SPK_E_NOMEM, "offset table allocation failed");
        offsets = offsets_tmp;
    }
    csr_finish_kernel<<<(unsigned)((m + 1 + 255) / 256), 256, 0, c->stream>>>(k1, e1buf, d_edges, m, bits, offsets, nn, idst, ival, ieid);
    weight_lists_kernel<<<gm, 256, 0, c->stream>>>(d_edges, we1, m, wdst, wval);
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AR1], c->stream));
    if (nparts == 0) nparts = 1;
    const unsigned long long e0 = (unsigned long long)n * part / nparts, e1 = (unsigned long long)n * (part + 1) / nparts;
    const size_t bm_bytes = (size_t)((nn + 31) / 32) * 4;
    if (bm_bytes <= 160 * 1024) {
        if (bm_bytes > 48 * 1024) SPK_CUDA(c, cudaFuncSetAttribute(aracne_bitmap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bm_bytes));
        const unsigned per_sm = (unsigned)std::max<size_t>(1, std::min<size_t>(8, (200 * 1024) / (bm_bytes + 1024)));
        unsigned long long blocks = (unsigned long long)c->sm_count * per_sm;
        if (blocks > nn) blocks = nn;
        aracne_bitmap_kernel<<<(unsigned)blocks, 256, bm_bytes, c->stream>>>(e0, e1, idst, ival, ieid, wdst, wval, offsets, nn, d_scan + 1, eps, d_flags);
    } else {
        unsigned long long blocks = (e1 - e0 + 7) / 8;
        const unsigned long long cap = (unsigned long long)c->sm_count * 32;
        if (blocks > cap) blocks = cap;
        if (blocks == 0) blocks = 1;
        aracne_search_kernel<<<(unsigned)blocks, 256, 0, c->stream>>>(d_edges, e0, e1, idst, ival, offsets, eps, d_flags);
    }
    c->tm.kernel_launches += 5 + (2 * bits + 7) / 8 + (bits + 7) / 8;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_AR2], c->stream));
    c->tm_pending |= TMP_ARB | TMP_AR;
    if (offsets_tmp) { SPK_CUDA(c, cudaStreamSynchronize(c->stream)); cudaFree(offsets_tmp); }
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

} // namespace spk
