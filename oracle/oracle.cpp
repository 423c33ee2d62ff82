// oracle.cpp -- CPU restatement of SpydrPick's hot path (TEST INFRASTRUCTURE ONLY).
//
// This file is the parity ORACLE and the timed CPU baseline. It is NOT part of the
// product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference legs may load it. The product path (spydrpick_b200/csrc) never links it.
//
// PINNED AGAINST THE REAL REFERENCE. The reference (santeripuranen/SpydrPick @ f6bc110, apegrunt @ 5c61810) ships no tests,
// golden vectors or fixtures of its own, and its build system needs Boost + legacy TBB (not installed). Its sources however
// compile IN PLACE against the small Boost / TBB stand-ins of oracle/ref_shims (oracle/Makefile.ref -> oracle/_ref/SpydrPick_ref,
// the reference program with its own main(); oracle/_ref/libspkref.so, C entry points over the same templates). This file is
// checked against that compiled reference: tests/test_ref_pins.py (normalize_and_get_mi_single bit-exact on > 1e5 tables,
// calculate_weights integer-exact incl. both quirks, the scan / MI-without-gaps / fences to 1e-14, Graph::sort, run_ARACNE and
// both writers byte-equal) and against the listings the reference program wrote for 16 runs (tests/golden/ref,
// tests/test_ref_golden.py). Known, documented deviation: the reference's per-pair path MI_solver::single (threshold
// estimator only) is defective for ~2 % of the pairs -- it disagrees with the reference's own scan; orc_mi_pairs follows the scan.
//
// Every function cites the reference file:line it restates. "ref:" paths are relative to
// /root/reference; "apegrunt/" = externals/apegrunt/include/apegrunt/, "misc/" =
// externals/apegrunt/include/misc/, "graph/" = externals/apegrunt/include/graph/.
//
// Data model: dense states uint8[n][L] row-major (sample-major), values 0..4 =
// a,t,c,g,GAP (ref: externals/apegrunt/src/StateVector_state_types.cpp:33-44).

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <limits>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <unordered_set>
#include <vector>

namespace {

constexpr int NS = 5;          // number of states incl. gap   (ref: apegrunt/StateVector_state_types.hpp:83-90)
constexpr int GAP = 4;
constexpr size_t BS = 32;      // StateBlock_size              (ref: apegrunt/Alignment_forward.h:29)

struct Edge { uint32_t i, j; double mi; };

int resolve_threads(int threads) {
    if (threads > 0) return threads;
    unsigned hc = std::thread::hardware_concurrency();
    return hc ? int(hc) : 1;
}

// simple static-chunk / dynamic-index parallel for
template <class F>
void parallel_for(size_t begin, size_t end, int threads, F f) {
    threads = resolve_threads(threads);
    if (threads <= 1 || end - begin <= 1) { for (size_t i = begin; i < end; ++i) f(i, 0); return; }
    std::atomic<size_t> next(begin);
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
        pool.emplace_back([&, t]() {
            for (;;) { size_t i = next.fetch_add(1); if (i >= end) break; f(i, t); }
        });
    }
    for (auto& th : pool) th.join();
}

// ---------------------------------------------------------------------------------------
// MI formula. ref: include/mi.hpp:146-181 (normalize_and_get_mi_single), helpers
// misc/Vector_operations.hpp:33-48 (sum, mask_sum), :66-71 (xlogx), :101-108 (mask_sum_xlogx),
// misc/Vector_interface.hpp:39 (masked constant vector).
// buffer is the 5x5 table in the reference's layout buffer[jstate*5 + istate]
// (misc/Crosstable.hpp:62-67) and is MODIFIED IN PLACE exactly as the reference does.
// ---------------------------------------------------------------------------------------
double normalize_and_get_mi_single(double* buffer, unsigned ipresence, unsigned jpresence, double pseudocount) {
    double masked_pc[NS];
    for (int i = 0; i < NS; ++i) masked_pc[i] = (ipresence & (1u << i)) ? pseudocount : 0.0;

    double normconst = 0.0;
    for (int j = 0; j < NS; ++j) {
        if (jpresence & (1u << j)) {
            double* row = buffer + j * NS;
            for (int i = 0; i < NS; ++i) row[i] += masked_pc[i];
            double s = 0.0;
            for (int i = 0; i < NS; ++i) if (ipresence & (1u << i)) s += row[i];
            normconst += s;
        }
    }
    double jointH = 0.0, icondH = 0.0, jcondH = 0.0;
    double jcondHvec[NS] = {0, 0, 0, 0, 0};
    for (int j = 0; j < NS; ++j) {
        if (jpresence & (1u << j)) {
            double* row = buffer + j * NS;
            for (int i = 0; i < NS; ++i) row[i] /= normconst;
            double s = 0.0;
            for (int i = 0; i < NS; ++i) if (ipresence & (1u << i)) s += row[i] * std::log(row[i]);
            jointH += s;
            double rs = 0.0;
            for (int i = 0; i < NS; ++i) rs += row[i];          // sum over ALL five states (mi.hpp:173)
            icondH += rs * std::log(rs);
            for (int i = 0; i < NS; ++i) jcondHvec[i] += row[i];
        }
    }
    for (int i = 0; i < NS; ++i) if (ipresence & (1u << i)) jcondH += jcondHvec[i] * std::log(jcondHvec[i]);
    return jointH - icondH - jcondH;
}

// weighted 5x5 crosstable of columns (ci, cj). ref: misc/Crosstable.hpp:62-67 (accumulate:
// data[jstate*N + istate] += weight), :239-263 (block variant, same orientation).
// The reference sums weights per unique-pattern intersection first; here in sample order
// (difference is fp64 re-association only, ~1e-16 relative).
void crosstable(const uint8_t* states, size_t n, size_t L, const double* w, size_t ci, size_t cj, double* buf) {
    for (int k = 0; k < NS * NS; ++k) buf[k] = 0.0;
    for (size_t s = 0; s < n; ++s) {
        unsigned a = states[s * L + ci], b = states[s * L + cj];
        buf[b * NS + a] += w[s];
    }
}

// per-column 5-bit state presence masks. ref: apegrunt/Alignment_impl_block_compressed_storage.hpp:1089-1128
void state_masks(const uint8_t* states, size_t n, size_t L, uint8_t* masks) {
    std::memset(masks, 0, L);
    for (size_t s = 0; s < n; ++s) {
        const uint8_t* row = states + s * L;
        for (size_t c = 0; c < L; ++c) masks[c] |= uint8_t(1u << row[c]);
    }
}

// ref: apegrunt/Distance.hpp:28-37
inline uint64_t linear_distance(uint64_t p1, uint64_t p2) {
    int64_t d = int64_t(p1) - int64_t(p2);
    return uint64_t(d < 0 ? -d : d);
}
inline uint64_t circular_distance(uint64_t p1, uint64_t p2, uint64_t circle) {
    uint64_t d = linear_distance(p1, p2);
    return std::min<uint64_t>(circle - d, d);   // size_t wrap when d > circle is intentional (reference behaviour)
}

// ref: graph/Graph.hpp:94-101 (id = node2<<32 | node1 with node1=min, node2=max), :235 (sort order)
inline uint64_t edge_id(const Edge& e) {
    uint32_t a = std::min(e.i, e.j), b = std::max(e.i, e.j);
    return (uint64_t(b) << 32) | uint64_t(a);
}
inline bool edge_less(const Edge& a, const Edge& b) {
    return (a.mi == b.mi) ? (edge_id(a) < edge_id(b)) : (a.mi > b.mi);
}

} // namespace

extern "C" {

// ---------------------------------------------------------------------------------------
// Sample-neighbour counts (int_weights) and weights.
// ref: apegrunt/Alignment_StateVector_weights.hpp:192-353 (calculate_weights),
//      apegrunt/StateVector_impl_block_compressed_alignment_storage.hpp:607-660 (is_similar_to_impl),
//      apegrunt/State_block.hpp:347-430 (count_identical = byte equality incl. gap==gap).
// counts[i] = 1 + #{j != i : similar(i,j)}   (multiplicity is always 1 in SpydrPick).
// similar(i,j):
//   L % 32 != 0 : ident(i,j) >= min_identical                       (exact)
//   L % 32 == 0 : ident(i,j) + ident_lastblock(i,j) >= min_identical when lastblock_quirk != 0
//                 (the reference's post-loop code counts the final block twice), else exact.
// ---------------------------------------------------------------------------------------
// `quirks` is a bit mask (include/spydrpick_b200.h SPK_QUIRK_*): bit 0 = the L % 32 == 0 final-block double count above;
// bit 1 = the SAMPLE-0 double count: calculate_weights first compares sample 0 with EVERY sample j in [0, n) -- itself
// included -- in its cache-priming pass (range_primer, Alignment_StateVector_weights.hpp:218-238, primer_ranges :275-283 cover
// [0, n)), and then compares all pairs i in [1, n), j in [0, i) -- sample 0 again -- in range_worker (:240-262, :287-295).
// So int_weights[0] = 1 + 2 (the self pair credits both ends) + 2 * #{j > 0 similar to 0}, and every j > 0 similar to
// sample 0 carries one extra count. Pinned against the compiled reference (tests/test_ref_pins.py).
int orc_neighbour_counts(const uint8_t* states, uint64_t n, uint64_t L, uint64_t min_identical,
                         int quirks, int threads, uint64_t* counts) {
    std::vector<std::atomic<uint64_t>> acc(n);
    for (auto& a : acc) a.store(1);
    const bool quirk = (quirks & 1) && L > 0 && (L % BS == 0);
    const size_t last0 = L >= BS ? L - BS : 0;
    auto similar = [&](size_t i, size_t j) {
        const uint8_t* ri = states + i * L; const uint8_t* rj = states + j * L;
        uint64_t eff = 0;
        for (size_t c = 0; c < L; ++c) eff += (ri[c] == rj[c]);
        if (quirk) { for (size_t c = last0; c < L; ++c) eff += (ri[c] == rj[c]); }
        return eff >= min_identical;
    };
    parallel_for(1, n, threads, [&](size_t i, int) {
        for (size_t j = 0; j < i; ++j)
            if (similar(i, j)) { acc[i].fetch_add(1); acc[j].fetch_add(1); }
    });
    if ((quirks & 2) && n > 0)
        for (size_t j = 0; j < n; ++j)
            if (similar(0, j)) { acc[0].fetch_add(1); acc[j].fetch_add(1); }
    for (size_t i = 0; i < n; ++i) counts[i] = acc[i].load();
    return 0;
}

// weights[i] = 1/counts[i]; ref: apegrunt/Alignment_StateVector_weights.hpp:349
// optional rescale (w * n) / sum(w): ref apegrunt/Alignment_utility.hpp:215,245
int orc_weights_from_counts(const uint64_t* counts, uint64_t n, int rescale, double* w) {
    for (uint64_t i = 0; i < n; ++i) w[i] = 1.0 / double(counts[i]);
    if (rescale) {
        double s = 0; for (uint64_t i = 0; i < n; ++i) s += w[i];
        for (uint64_t i = 0; i < n; ++i) w[i] = w[i] * double(n) / s;      // (w * n) / n_eff, left to right like the reference
    }
    return 0;
}

int orc_state_masks(const uint8_t* states, uint64_t n, uint64_t L, uint8_t* masks) {
    state_masks(states, n, L, masks);
    return 0;
}

// MI from an explicit 5x5 table (buffer[jstate*5+istate]); modifies buffer (reference semantics).
double orc_mi_from_table(double* buffer, unsigned ipresence, unsigned jpresence, double pseudocount) {
    return normalize_and_get_mi_single(buffer, ipresence, jpresence, pseudocount);
}

// MI of one column pair with full presence masks; ref: include/mi.hpp:183-224 (single, cached=false)
// if mi_wo_gaps != NULL also returns the reference's literal "MI without gaps" (second pass on the
// already-normalised buffer with gap bits removed from both masks; ref include/mi.hpp:108-135,465-489).
double orc_mi_pair(const uint8_t* states, uint64_t n, uint64_t L, const double* w, const uint8_t* masks,
                   uint64_t ci, uint64_t cj, double pseudocount, double* mi_wo_gaps) {
    double buf[NS * NS];
    crosstable(states, n, L, w, ci, cj, buf);
    double mi = normalize_and_get_mi_single(buf, masks[ci], masks[cj], pseudocount);
    if (mi_wo_gaps) {
        const unsigned nogap = ~(1u << GAP);
        *mi_wo_gaps = normalize_and_get_mi_single(buf, masks[ci] & nogap, masks[cj] & nogap, pseudocount);
    }
    return mi;
}

// MI for an explicit pair list (threshold estimation); ref: include/SpydrPick.hpp:210-255
int orc_mi_pairs(const uint8_t* states, uint64_t n, uint64_t L, const double* w,
                 const uint32_t* pi, const uint32_t* pj, uint64_t npairs, double pseudocount,
                 int threads, double* mi) {
    std::vector<uint8_t> masks(L);
    state_masks(states, n, L, masks.data());
    // column-major copy for cache friendliness
    std::vector<uint8_t> colmaj(size_t(n) * L);
    for (size_t s = 0; s < n; ++s) for (size_t c = 0; c < L; ++c) colmaj[c * n + s] = states[s * L + c];
    parallel_for(0, npairs, threads, [&](size_t p, int) {
        double buf[NS * NS] = {0};
        const uint8_t* a = colmaj.data() + size_t(pi[p]) * n;
        const uint8_t* b = colmaj.data() + size_t(pj[p]) * n;
        for (size_t s = 0; s < n; ++s) buf[b[s] * NS + a[s]] += w[s];
        mi[p] = normalize_and_get_mi_single(buf, masks[pi[p]], masks[pj[p]], pseudocount);
    });
    return 0;
}

// ---------------------------------------------------------------------------------------
// Full upper-triangle scan. ref: include/mi.hpp:377-508 (MI_solver::operator()),
// include/SpydrPick.hpp:132-168 (get_MI_network).
//  * every pair i<j: mi = MI(i,j)
//  * if distance(pos[i],pos[j]) > ld_threshold: colmax[i]=max(colmax[i],mi), same for j
//    distance = linear or circular with circle = circle_size (reference passes L, the
//    FILTERED column count: apegrunt/Distance.hpp:67-71)
//  * if threshold < mi (strict): edge saved; if gap present in column i or j the
//    MI-without-gaps (literal second pass) is saved to the wo_gaps list.
// Results are returned in malloc'ed arrays (free with orc_free). Edge order is
// unspecified (the reference's order depends on TBB scheduling); callers sort.
// ---------------------------------------------------------------------------------------
struct orc_scan_result {
    Edge* edges; uint64_t n_edges;
    Edge* edges_wo_gaps; uint64_t n_wo_gaps;
    double* colmax;            // [L], lowest double if the column had no admissible pair
};

int orc_mi_scan(const uint8_t* states, uint64_t n, uint64_t L, const double* w, const uint64_t* pos,
                double threshold, double pseudocount, uint64_t ld_threshold, int linear_genome,
                uint64_t circle_size, int threads, orc_scan_result* out) {
    threads = resolve_threads(threads);
    std::vector<uint8_t> masks(L);
    state_masks(states, n, L, masks.data());
    std::vector<uint8_t> colmaj(size_t(n) * L);
    for (size_t s = 0; s < n; ++s) for (size_t c = 0; c < L; ++c) colmaj[c * n + s] = states[s * L + c];

    std::vector<std::vector<Edge>> tl_edges(threads), tl_wo(threads);
    std::vector<std::vector<double>> tl_colmax(threads, std::vector<double>(L, std::numeric_limits<double>::lowest()));
    const unsigned nogap = ~(1u << GAP);

    parallel_for(0, L, threads, [&](size_t i, int t) {
        auto& edges = tl_edges[t]; auto& wo = tl_wo[t]; auto& cm = tl_colmax[t];
        const uint8_t* a = colmaj.data() + i * n;
        for (size_t j = i + 1; j < L; ++j) {
            const uint8_t* b = colmaj.data() + j * n;
            double buf[NS * NS] = {0};
            for (size_t s = 0; s < n; ++s) buf[b[s] * NS + a[s]] += w[s];
            const double mi = normalize_and_get_mi_single(buf, masks[i], masks[j], pseudocount);
            const uint64_t d = linear_genome ? linear_distance(pos[i], pos[j]) : circular_distance(pos[i], pos[j], circle_size);
            if (d > ld_threshold) { if (mi > cm[i]) cm[i] = mi; if (mi > cm[j]) cm[j] = mi; }
            if (threshold < mi) {
                edges.push_back(Edge{uint32_t(i), uint32_t(j), mi});
                if ((masks[i] | masks[j]) & (1u << GAP)) {
                    const double mi2 = normalize_and_get_mi_single(buf, masks[i] & nogap, masks[j] & nogap, pseudocount);
                    wo.push_back(Edge{uint32_t(i), uint32_t(j), mi2});
                }
            }
        }
    });

    size_t ne = 0, nw = 0;
    for (int t = 0; t < threads; ++t) { ne += tl_edges[t].size(); nw += tl_wo[t].size(); }
    out->edges = (Edge*)std::malloc(std::max<size_t>(ne, 1) * sizeof(Edge));
    out->edges_wo_gaps = (Edge*)std::malloc(std::max<size_t>(nw, 1) * sizeof(Edge));
    out->colmax = (double*)std::malloc(std::max<size_t>(L, 1) * sizeof(double));
    size_t pe = 0, pw = 0;
    for (int t = 0; t < threads; ++t) {
        std::copy(tl_edges[t].begin(), tl_edges[t].end(), out->edges + pe); pe += tl_edges[t].size();
        std::copy(tl_wo[t].begin(), tl_wo[t].end(), out->edges_wo_gaps + pw); pw += tl_wo[t].size();
    }
    out->n_edges = ne; out->n_wo_gaps = nw;
    for (size_t c = 0; c < L; ++c) {
        double m = std::numeric_limits<double>::lowest();
        for (int t = 0; t < threads; ++t) m = std::max(m, tl_colmax[t][c]);   // join: mi.hpp:355-361
        out->colmax[c] = m;
    }
    return 0;
}

void orc_free_scan(orc_scan_result* r) {
    std::free(r->edges); std::free(r->edges_wo_gaps); std::free(r->colmax);
    r->edges = r->edges_wo_gaps = nullptr; r->colmax = nullptr;
}

// Tukey fences from the per-column maxima. ref: include/mi.hpp:267-281 (quartile index npos/4*Q,
// integer division first), include/SpydrPick.hpp:148-152.
int orc_outlier_thresholds(const double* colmax, uint64_t L, double* outlier, double* extreme) {
    std::vector<double> v(colmax, colmax + L);
    std::sort(v.begin(), v.end());
    const double Q1 = v[L / 4 * 1], Q3 = v[L / 4 * 3];
    *outlier = Q3 + 1.5 * (Q3 - Q1);
    *extreme = Q3 + 3.0 * (Q3 - Q1);
    return 0;
}

// ref: graph/Graph.hpp:234-242
int orc_sort_edges(Edge* edges, uint64_t n_edges) {
    std::sort(edges, edges + n_edges, edge_less);
    return 0;
}

// ---------------------------------------------------------------------------------------
// ARACNE, literal block-streamed restatement. ref: include/ARACNE.hpp:51-88 (remap),
// :91-150 (block_reader), :190-244 (block_sorter), :246-368 (block_processor), :371-405
// (result_storer), :411-548 (aracne). Input edges must be sorted (weight desc, id asc).
// direct_flags[e] = 1 for DIRECT edges (reference: bool(edge) after run_ARACNE, printed in
// the ARACNE column), 0 for edges marked indirect.
// The threshold==0 "block_start_fixed" rewind (:482-487) indexes indirect_edges out of
// bounds in the reference (UB); here the container is sized for the rewound range, which is
// the evident intent.
// ---------------------------------------------------------------------------------------
int orc_aracne_blocked(const Edge* edges, uint64_t n_edges_in, double threshold, uint32_t block_size,
                       uint8_t* direct_flags) {
    const uint32_t n_edges = uint32_t(n_edges_in);              // ref :416 (uint32_t)
    if (block_size == 0) block_size = 16384;
    // remap_and_initialize: dense ids in first-appearance order
    std::vector<uint32_t> r1(n_edges), r2(n_edges);
    std::vector<std::vector<std::pair<uint32_t, uint64_t>>> nbh;
    {
        std::vector<int64_t> map;   // node id -> dense id
        auto remap = [&](uint32_t v) -> uint32_t {
            if (v >= map.size()) map.resize(size_t(v) + 1, -1);
            if (map[v] < 0) { map[v] = int64_t(nbh.size()); nbh.emplace_back(); }
            return uint32_t(map[v]);
        };
        for (uint32_t e = 0; e < n_edges; ++e) {
            uint32_t a = remap(std::min(edges[e].i, edges[e].j));
            uint32_t b = remap(std::max(edges[e].i, edges[e].j));
            r1[e] = std::min(a, b); r2[e] = std::max(a, b);     // Edge ctor orders (min,max): graph/Graph.hpp:52-58
        }
    }
    std::vector<uint8_t> indirect(n_edges, 0);
    auto bsearch = [&](const std::vector<std::pair<uint32_t, uint64_t>>& v, uint32_t key) -> uint64_t {
        int32_t a = 0, b = int32_t(v.size()) - 1;
        while (a <= b) {
            int32_t mid = (a + b) / 2;
            if (v[mid].first == key) return v[mid].second;
            if (v[mid].first > key) b = mid - 1; else a = mid + 1;
        }
        return std::numeric_limits<uint64_t>::max();
    };
    for (size_t block_start = 0; block_start < n_edges; block_start += block_size) {
        const size_t block_end = std::min<size_t>(block_start + block_size, n_edges);
        // block_reader + block_sorter
        std::vector<uint32_t> touched;
        for (size_t e = block_start; e < block_end; ++e) {
            nbh[r1[e]].emplace_back(r2[e], e); nbh[r2[e]].emplace_back(r1[e], e);
            touched.push_back(r1[e]); touched.push_back(r2[e]);
        }
        std::sort(touched.begin(), touched.end());
        touched.erase(std::unique(touched.begin(), touched.end()), touched.end());
        for (uint32_t v : touched) std::sort(nbh[v].begin(), nbh[v].end());
        // block_processor
        size_t block_start_fixed = block_start;
        if (threshold == 0.0 && block_start > 0) {
            while (block_start_fixed > 0 && edges[block_start_fixed - 1].mi == edges[block_start].mi) --block_start_fixed;
        }
        std::vector<size_t> to_set;
        for (size_t e = block_start_fixed; e < block_end; ++e) {
            const uint32_t node1 = r1[e], node2 = r2[e];
            const double mi1 = edges[e].mi;
            for (const auto& nb : nbh[node1]) {
                if (nb.first == node2) continue;
                if (nb.second >= block_start_fixed && nb.second < e) continue;
                const uint64_t res = bsearch(nbh[node2], nb.first);
                if (res == std::numeric_limits<uint64_t>::max()) continue;
                const double mi2 = edges[nb.second].mi, mi3 = edges[res].mi;
                const double midval = std::max(std::min(mi1, mi2), std::min(std::max(mi1, mi2), mi3));
                const double minval = std::min(mi1, std::min(mi2, mi3));
                if (midval - minval >= threshold) {
                    if (mi1 == minval) to_set.push_back(e);
                    if (mi2 == minval) to_set.push_back(nb.second);
                    if (mi3 == minval) to_set.push_back(res);
                }
            }
        }
        for (size_t e : to_set) indirect[e] = 1;
    }
    for (uint32_t e = 0; e < n_edges; ++e) direct_flags[e] = indirect[e] ? 0 : 1;   // "Booleans swapped" :399-405
    return 0;
}

// ARACNE closed form: edge e=(u,v) is indirect iff there is a common neighbour k
// with w_e == min(w_e,w_uk,w_vk) and median - min >= threshold. Order independent; used to
// cross-check the literal version and as the spec of the CUDA kernel.
int orc_aracne_closed(const Edge* edges, uint64_t n_edges, double threshold, int threads, uint8_t* direct_flags) {
    uint32_t nn = 0;
    for (uint64_t e = 0; e < n_edges; ++e) nn = std::max(nn, std::max(edges[e].i, edges[e].j) + 1);
    std::vector<std::vector<std::pair<uint32_t, double>>> nbh(nn);
    for (uint64_t e = 0; e < n_edges; ++e) {
        nbh[edges[e].i].emplace_back(edges[e].j, edges[e].mi);
        nbh[edges[e].j].emplace_back(edges[e].i, edges[e].mi);
    }
    for (auto& v : nbh) std::sort(v.begin(), v.end());
    parallel_for(0, n_edges, threads, [&](size_t e, int) {
        const auto& A = nbh[edges[e].i]; const auto& B = nbh[edges[e].j];
        const double w = edges[e].mi;
        bool ind = false;
        size_t a = 0, b = 0;
        while (a < A.size() && b < B.size() && !ind) {
            if (A[a].first < B[b].first) ++a;
            else if (A[a].first > B[b].first) ++b;
            else {
                const double w2 = A[a].second, w3 = B[b].second;
                const double midval = std::max(std::min(w, w2), std::min(std::max(w, w2), w3));
                const double minval = std::min(w, std::min(w2, w3));
                if (midval - minval >= threshold && w == minval) ind = true;
                ++a; ++b;
            }
        }
        direct_flags[e] = ind ? 0 : 1;
    });
    return 0;
}

// ---------------------------------------------------------------------------------------
// MI threshold estimation. ref: include/SpydrPick.hpp:258-282 (determine_threshold_pairs),
// :171-207 (sample_pairs; the reference seeds mt19937 from steady_clock -- here an explicit
// seed), :285-343 (determine_MI_threshold).
// ---------------------------------------------------------------------------------------
uint64_t orc_determine_threshold_pairs(uint64_t threshold_pairs, uint64_t possible_pairs, double threshold_percentile) {
    if (threshold_pairs == 0) {
        threshold_pairs = 100000;
        const uint64_t desired_idx_from_end = 100, desired_max = 500000;
        while (double(threshold_pairs) - threshold_percentile * double(threshold_pairs) < double(desired_idx_from_end) &&
               threshold_pairs < desired_max) threshold_pairs += 10000;
    }
    if (possible_pairs / 10 < threshold_pairs) threshold_pairs = possible_pairs / 10;
    return threshold_pairs;
}

// fills pi/pj with `count` unique pairs i<j, indices in [0,max_range]
int orc_sample_pairs(uint64_t count, uint64_t max_range, uint64_t seed, uint32_t* pi, uint32_t* pj) {
    std::mt19937 gen((uint32_t)seed);
    std::uniform_int_distribution<uint64_t> dist(0, max_range);
    std::unordered_set<uint64_t> seen;
    std::vector<uint64_t> order;
    while (seen.size() < count) {
        uint64_t a = dist(gen), b = dist(gen);
        while (a == b) { a = dist(gen); b = dist(gen); }
        if (a > b) std::swap(a, b);
        if (seen.insert((a << 32) | b).second) order.push_back((a << 32) | b);
    }
    for (uint64_t k = 0; k < count; ++k) { pi[k] = uint32_t(order[k] >> 32); pj[k] = uint32_t(order[k]); }
    return 0;
}

double orc_determine_mi_threshold(const uint8_t* states, uint64_t n, uint64_t L, const double* w,
                                  uint64_t n_values, uint64_t threshold_pairs_opt, uint64_t iterations,
                                  double pseudocount, uint64_t seed, int threads) {
    const uint64_t possible_pairs = L * (L - 1) / 2;
    const double pct = 1.0 - double(n_values) / double(possible_pairs);
    const uint64_t tp = orc_determine_threshold_pairs(threshold_pairs_opt, possible_pairs, pct);
    const uint64_t idx = uint64_t(pct * double(tp));
    std::vector<double> thresholds;
    std::vector<uint32_t> pi(tp), pj(tp);
    std::vector<double> mi(tp);
    for (uint64_t it = 0; it < iterations; ++it) {
        orc_sample_pairs(tp, L - 1, seed + it, pi.data(), pj.data());
        orc_mi_pairs(states, n, L, w, pi.data(), pj.data(), tp, pseudocount, threads, mi.data());
        std::nth_element(mi.begin(), mi.begin() + idx, mi.end());
        thresholds.push_back(mi[idx]);
    }
    const size_t median_idx = thresholds.size() / 2 - (thresholds.size() % 2 ? 0 : 1);
    std::nth_element(thresholds.begin(), thresholds.begin() + median_idx, thresholds.end());
    return thresholds[median_idx];
}

// ---------------------------------------------------------------------------------------
// Output writers (text, byte-exact). ref: graph/Graph_output_format.hpp:49-74 (edges file),
// include/SpydrPick.hpp:89-129 (outliers file). Distances are computed on the OUTPUT
// indices (translated position + base) with circle = n_original_positions
// (graph/Graph_output_format.hpp:35, include/SpydrPick.hpp:75).
// Both return a malloc'ed NUL-terminated string (free with orc_free_str).
// ---------------------------------------------------------------------------------------
static void append_fixed(std::string& s, double v, int prec) {
    char tmp[64]; std::snprintf(tmp, sizeof tmp, "%.*f", prec, v); s += tmp;
}

char* orc_format_edges(const Edge* edges, const uint8_t* direct_flags, uint64_t n_edges, const uint64_t* pos,
                       uint64_t base_index, int linear_genome, uint64_t n_original_positions) {
    std::string s;
    for (uint64_t e = 0; e < n_edges; ++e) {
        const uint32_t n1 = std::min(edges[e].i, edges[e].j), n2 = std::max(edges[e].i, edges[e].j);
        const uint64_t i1 = pos[n1] + base_index, i2 = pos[n2] + base_index;
        const uint64_t d = linear_genome ? linear_distance(i1, i2) : circular_distance(i1, i2, n_original_positions);
        s += std::to_string(i1); s += ' '; s += std::to_string(i2); s += ' '; s += std::to_string(d); s += ' ';
        s += (direct_flags && direct_flags[e]) ? '1' : '0'; s += ' ';
        append_fixed(s, edges[e].mi, 6); s += '\n';
    }
    char* out = (char*)std::malloc(s.size() + 1);
    std::memcpy(out, s.c_str(), s.size() + 1);
    return out;
}

// edges: sorted desc; wo_gaps: any order (sorted by id here, as the formatter's ctor does)
char* orc_format_outliers(const Edge* edges, const uint8_t* direct_flags, uint64_t n_edges,
                          const Edge* wo_gaps, uint64_t n_wo, const uint64_t* pos, uint64_t base_index,
                          int linear_genome, uint64_t n_original_positions, uint64_t ld_threshold,
                          double outlier_threshold, double extreme_threshold) {
    std::vector<Edge> wo(wo_gaps, wo_gaps + n_wo);
    std::sort(wo.begin(), wo.end(), [](const Edge& a, const Edge& b) { return edge_id(a) < edge_id(b); });
    std::string s;
    for (uint64_t e = 0; e < n_edges; ++e) {
        const Edge& ed = edges[e];
        if (ed.mi < outlier_threshold) break;
        auto it = std::lower_bound(wo.begin(), wo.end(), ed, [](const Edge& a, const Edge& b) { return edge_id(a) < edge_id(b); });
        const double w2 = (it != wo.end() && edge_id(*it) == edge_id(ed)) ? it->mi : ed.mi;
        const uint32_t n1 = std::min(ed.i, ed.j), n2 = std::max(ed.i, ed.j);
        const uint64_t i1 = pos[n1] + base_index, i2 = pos[n2] + base_index;
        const uint64_t d = linear_genome ? linear_distance(i1, i2) : circular_distance(i1, i2, n_original_positions);
        if (d > ld_threshold) {
            s += std::to_string(i1); s += ' '; s += std::to_string(i2); s += ' '; s += std::to_string(d); s += ' ';
            s += (direct_flags && direct_flags[e]) ? '1' : '0'; s += ' ';
            append_fixed(s, ed.mi, 6); s += ' ';
            append_fixed(s, w2, 6); s += ' ';
            append_fixed(s, (1.0 - (w2 / ed.mi)) * 100, 1); s += ' ';
            s += (ed.mi > extreme_threshold) ? '1' : '0'; s += '\n';
        }
    }
    char* out = (char*)std::malloc(s.size() + 1);
    std::memcpy(out, s.c_str(), s.size() + 1);
    return out;
}

void orc_free_str(char* p) { std::free(p); }

// ---------------------------------------------------------------------------------------
// Column filter. ref: apegrunt/Alignment_filter.hpp:284-318: unweighted per-column state
// frequencies; keep a column iff (#non-gap states with frequency >= maf) > 1 and gap
// frequency <= gap_threshold. keep[c] = 1/0; returns number kept.
// ---------------------------------------------------------------------------------------
uint64_t orc_filter_columns(const uint8_t* states, uint64_t n, uint64_t L, double maf, double gap_threshold, uint8_t* keep) {
    std::vector<uint64_t> cnt(L * NS, 0);
    for (uint64_t s = 0; s < n; ++s) for (uint64_t c = 0; c < L; ++c) cnt[c * NS + states[s * L + c]]++;
    uint64_t kept = 0;
    for (uint64_t c = 0; c < L; ++c) {
        int nstates = 0;
        for (int a = 0; a < GAP; ++a) if (double(cnt[c * NS + a]) / double(n) >= maf) ++nstates;
        const double gf = double(cnt[c * NS + GAP]) / double(n);
        keep[c] = (nstates > 1 && gf <= gap_threshold) ? 1 : 0;
        kept += keep[c];
    }
    return kept;
}

int orc_hardware_threads(void) { return resolve_threads(0); }

} // extern "C"
