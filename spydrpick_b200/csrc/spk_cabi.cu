// spk_cabi.cu -- the C-ABI (include/spydrpick_b200.h): context, buffers, stage orchestration.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdarg>
#include <cstring>
#include <random>
#include <unordered_set>
#include "spk_internal.h"

namespace spk {

static thread_local std::string g_create_error;

int fail(Ctx* c, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    if (c) c->err = buf; else g_create_error = buf;
    if (c && c->verbose) fprintf(stderr, "spydrpick_b200: error %d: %s\n", code, buf);
    return code;
}
int cuda_check(Ctx* c, cudaError_t e, const char* what) {
    if (e == cudaSuccess) return SPK_OK;
    return fail(c, SPK_E_CUDA, "%s: %s", what, cudaGetErrorString(e));
}
void* scratch(Ctx* c, size_t bytes) {
    if (bytes <= c->scratch_bytes) return c->d_scratch;
    cudaStreamSynchronize(c->stream);
    if (c->d_scratch) cudaFree(c->d_scratch);
    c->d_scratch = nullptr; c->scratch_bytes = 0;
    const size_t want = bytes + bytes / 8;
    if (cudaMalloc(&c->d_scratch, want) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    c->scratch_bytes = want;
    return c->d_scratch;
}

template <class P> static void free_dev(P*& p) { if (p) { cudaFree(p); p = nullptr; } }

static void free_alignment(Ctx* c) {
    free_dev(c->d_states); free_dev(c->d_colmajor); free_dev(c->d_planes); free_dev(c->d_pos); free_dev(c->d_masks);
    free_dev(c->d_w); free_dev(c->d_marg); free_dev(c->d_colstat); free_dev(c->d_U); free_dev(c->d_V);
    free_dev(c->d_colmax); free_dev(c->d_colmax32); free_dev(c->d_counts);
    i8_free(c); mi_exact_free(c);
    c->operands_terms = 0; c->operands_cap_terms = 0; c->stats_valid = false;
}
static void free_scan(Ctx* c) {
    free_dev(c->d_cand); free_dev(c->d_cand2); c->cand2_cap = 0; free_dev(c->d_edges); free_dev(c->d_wo); free_dev(c->d_refined); free_dev(c->d_refined_wo);
    c->cand_cap = c->edges_cap = c->wo_cap = c->refined_cap = 0;
}

template <class P> static int dev_alloc(Ctx* c, P*& p, size_t bytes, const char* what) {
    free_dev(p);
    if (cudaMalloc(&p, bytes ? bytes : 16) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "device allocation failed: %s (%zu bytes)", what, bytes); }
    return SPK_OK;
}

__global__ void ordered_to_double_kernel(unsigned long long* v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const double d = f64_from_ordered(v[i]); v[i] = (unsigned long long)__double_as_longlong(d); }
}
__global__ void fill_u64_kernel(unsigned long long* p, unsigned long long v, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

} // namespace spk

using namespace spk;
struct spk_ctx : public spk::Ctx {};

extern "C" {

const char* spk_version(void) { return "spydrpick_b200 0.1 (sm_100a)"; }

int spk_create(spk_ctx** out, int device, int verbose) {
    if (!out) return SPK_E_INVALID;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, SPK_E_CUDA, "no CUDA device available (this library has no CPU fallback)");
    }
    if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) return fail(nullptr, SPK_E_CUDA, "cudaGetDevice failed"); }
    if (device >= ndev) return fail(nullptr, SPK_E_INVALID, "device %d out of range (%d devices)", device, ndev);
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail(nullptr, SPK_E_CUDA, "cudaGetDeviceProperties failed");
    if (prop.major != 10) return fail(nullptr, SPK_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor);
    spk_ctx* c = new spk_ctx();
    c->device = device; c->verbose = verbose; c->sm_count = prop.multiProcessorCount;
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c; return fail(nullptr, SPK_E_CUDA, "cannot create stream on device %d", device);
    }
    c->stream = c->own_stream;
    for (auto& e : c->ev) cudaEventCreate(&e);
    cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&c->ev_ready, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_dl, cudaEventDisableTiming);
    if (cudaMalloc(&c->d_counters, 8 * sizeof(unsigned long long)) != cudaSuccess || cudaMalloc(&c->d_tilemap, 1 << 20) != cudaSuccess) {
        delete c; return fail(nullptr, SPK_E_NOMEM, "context allocation failed");
    }
    *out = c;
    return SPK_OK;
}

void spk_destroy(spk_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    free_alignment(c); free_scan(c);
    free_dev(c->d_scratch); free_dev(c->d_counters);
    if (c->hp_edges) cudaFreeHost(c->hp_edges);
    if (c->hp_wo) cudaFreeHost(c->hp_wo);
    if (c->d_tilemap) { cudaFree(c->d_tilemap); c->d_tilemap = nullptr; }
    for (auto& e : c->ev) if (e) cudaEventDestroy(e);
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->ev_ready) cudaEventDestroy(c->ev_ready);
    if (c->ev_dl) cudaEventDestroy(c->ev_dl);
    if (c->d_flags_buf) cudaFree(c->d_flags_buf);
    if (c->h_mail) cudaFreeHost(c->h_mail);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

const char* spk_last_error(const spk_ctx* c) { return c ? c->err.c_str() : g_create_error.c_str(); }

int spk_set_stream(spk_ctx* c, void* s) {
    if (!c) return SPK_E_INVALID;
    cudaStreamSynchronize(c->stream);
    c->stream = s ? (cudaStream_t)s : c->own_stream;
    return SPK_OK;
}

static int set_alignment_impl(spk_ctx* c, const uint8_t* states, uint32_t n, uint32_t L, const uint64_t* pos, cudaMemcpyKind kind) {
    if (!c || !states || n == 0 || L < 2) return c ? fail(c, SPK_E_INVALID, "set_alignment: need states, n >= 1, L >= 2") : SPK_E_INVALID;
    if (L >= (1u << 29)) return fail(c, SPK_E_INVALID, "L too large");
    SPK_CUDA(c, cudaSetDevice(c->device));
    const bool same_shape = (c->n == n && c->L == L && c->d_states);
    if (!same_shape) {
        free_alignment(c);
        c->n = n; c->L = L;
        c->n_pad = (n + 127) / 128 * 128;                    // K blocks: 64 samples (bf16 engine), 128 (int8 engine)
        c->L_pad = (L + TILE_J - 1) / TILE_J * TILE_J;
        c->W32 = (L + 31) / 32;
        int r;
        if ((r = dev_alloc(c, c->d_states, (size_t)n * L, "states"))) return r;
        if ((r = dev_alloc(c, c->d_colmajor, (size_t)c->L_pad * c->n_pad, "colmajor"))) return r;
        if ((r = dev_alloc(c, c->d_planes, (size_t)3 * n * c->W32 * 4, "planes"))) return r;
        if ((r = dev_alloc(c, c->d_pos, (size_t)c->L_pad * 8, "pos"))) return r;
        if ((r = dev_alloc(c, c->d_masks, (size_t)c->L_pad, "masks"))) return r;
        if ((r = dev_alloc(c, c->d_w, (size_t)c->n_pad * 8, "weights"))) return r;
        if ((r = dev_alloc(c, c->d_marg, (size_t)c->L_pad * 5 * 8, "marginals"))) return r;
        if ((r = dev_alloc(c, c->d_colstat, (size_t)c->L_pad * sizeof(ColStat), "colstat"))) return r;
        if ((r = dev_alloc(c, c->d_colmax, (size_t)c->L_pad * 8, "colmax"))) return r;
        if ((r = dev_alloc(c, c->d_colmax32, (size_t)c->L_pad * 4, "colmax32"))) return r;
        if ((r = dev_alloc(c, c->d_counts, (size_t)n * 8, "counts"))) return r;
    }
    c->operands_terms = 0; c->stats_valid = false; c->host_edges_n = 0; c->g_state = 0; c->d_edges_sorted_n = 0;
    SPK_CUDA(c, cudaEventRecord(c->ev[0], c->stream));
    SPK_CUDA(c, cudaMemcpyAsync(c->d_states, states, (size_t)n * L, kind, c->stream));
    std::vector<uint64_t> hp(c->L_pad, 0);
    for (uint32_t k = 0; k < L; ++k) hp[k] = pos ? pos[k] : k;
    SPK_CUDA(c, cudaMemcpyAsync(c->d_pos, hp.data(), (size_t)c->L_pad * 8, cudaMemcpyHostToDevice, c->stream));
    std::vector<double> ones(c->n_pad, 0.0);
    for (uint32_t s = 0; s < n; ++s) ones[s] = 1.0;
    SPK_CUDA(c, cudaMemcpyAsync(c->d_w, ones.data(), (size_t)c->n_pad * 8, cudaMemcpyHostToDevice, c->stream));
    c->w_total = (double)n;
    c->h_w.assign(n, 1.0);
    SPK_CUDA(c, cudaEventRecord(c->ev[1], c->stream));
    int r = prep_alignment(c); if (r) return r;
    SPK_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
    SPK_CUDA(c, cudaEventSynchronize(c->ev[2]));
    r = i8_build_layout(c); if (r) return r;
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.upload_ms, c->ev[0], c->ev[1]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.prep_ms, c->ev[1], c->ev[2]));
    return SPK_OK;
}

int spk_set_alignment(spk_ctx* c, const uint8_t* states, uint32_t n, uint32_t L, const uint64_t* pos) {
    return set_alignment_impl(c, states, n, L, pos, cudaMemcpyHostToDevice);
}

// states already in this device's memory (e.g. row slices uploaded by the ranks of a box and all-gathered over NVLink)
int spk_set_alignment_device(spk_ctx* c, const void* d_states, uint32_t n, uint32_t L, const uint64_t* pos) {
    if (c && d_states) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, d_states) != cudaSuccess || at.type != cudaMemoryTypeDevice || at.device != c->device) {
            cudaGetLastError();
            return fail(c, SPK_E_INVALID, "set_alignment_device: d_states is not memory of device %d", c->device);
        }
    }
    return set_alignment_impl(c, (const uint8_t*)d_states, n, L, pos, cudaMemcpyDeviceToDevice);
}

// ---- column filter pre-pass (apegrunt/Alignment_filter.hpp:250-339) ----
namespace {
struct RawUpload {                                       // the raw alignment in HBM for the duration of one call
    uint8_t* d_raw = nullptr; uint32_t* d_counts = nullptr;
    ~RawUpload() { if (d_raw) cudaFree(d_raw); if (d_counts) cudaFree(d_counts); }
};
int upload_and_count(spk_ctx* c, const uint8_t* states, uint32_t n, uint64_t L, RawUpload& u, std::vector<uint32_t>& counts) {
    if (cudaMalloc(&u.d_raw, (size_t)n * L) != cudaSuccess || cudaMalloc(&u.d_counts, (size_t)L * NSTATES * 4) != cudaSuccess) {
        cudaGetLastError(); return fail(c, SPK_E_NOMEM, "column filter: cannot allocate %zu bytes for the raw alignment", (size_t)n * L);
    }
    SPK_CUDA(c, cudaMemcpyAsync(u.d_raw, states, (size_t)n * L, cudaMemcpyHostToDevice, c->stream));
    int r = column_state_counts_device(c, u.d_raw, n, L, u.d_counts); if (r) return r;
    counts.resize((size_t)L * NSTATES);
    SPK_CUDA(c, cudaMemcpyAsync(counts.data(), u.d_counts, (size_t)L * NSTATES * 4, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}
}  // namespace

int spk_column_state_counts(spk_ctx* c, const uint8_t* states, uint32_t n, uint64_t L, uint32_t* counts) {
    if (!c || !states || !counts || n == 0 || L == 0) return c ? fail(c, SPK_E_INVALID, "column_state_counts: need states, counts, n >= 1, L >= 1") : SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    RawUpload u; std::vector<uint32_t> h;
    int r = upload_and_count(c, states, n, L, u, h); if (r) return r;
    std::memcpy(counts, h.data(), h.size() * 4);
    return SPK_OK;
}

int spk_set_alignment_filtered(spk_ctx* c, const uint8_t* states, uint32_t n, uint64_t L, const uint64_t* pos,
                               double maf_threshold, double gap_threshold, uint8_t* keep, uint64_t* n_kept) {
    if (!c || !states || !keep || !n_kept || n == 0 || L < 2) return c ? fail(c, SPK_E_INVALID, "set_alignment_filtered: need states, keep, n_kept, n >= 1, L >= 2") : SPK_E_INVALID;
    if (L > 0xffffffffull) return fail(c, SPK_E_INVALID, "L too large");
    SPK_CUDA(c, cudaSetDevice(c->device));
    RawUpload u; std::vector<uint32_t> cnt;
    int r = upload_and_count(c, states, n, L, u, cnt); if (r) return r;
    // the decision is the reference's, in its arithmetic (frequencies as doubles; Alignment_filter.hpp:284-318)
    std::vector<uint32_t> idx; std::vector<uint64_t> kpos;
    for (uint64_t col = 0; col < L; ++col) {
        int nsig = 0;
        for (int a = 0; a < GAP; ++a) if (double(cnt[col * NSTATES + a]) / double(n) >= maf_threshold) ++nsig;
        const bool k = nsig > 1 && double(cnt[col * NSTATES + GAP]) / double(n) <= gap_threshold;
        keep[col] = k ? 1 : 0;
        if (k) { idx.push_back((uint32_t)col); kpos.push_back(pos ? pos[col] : col); }
    }
    *n_kept = idx.size();
    if (idx.size() < 2) return fail(c, SPK_E_INVALID, "fewer than 2 columns pass the MAF / gap filter");
    const uint32_t Lk = (uint32_t)idx.size();
    uint32_t* d_idx = nullptr; uint8_t* d_f = nullptr;
    if (cudaMalloc(&d_idx, (size_t)Lk * 4) != cudaSuccess || cudaMalloc(&d_f, (size_t)n * Lk) != cudaSuccess) {
        cudaGetLastError(); if (d_idx) cudaFree(d_idx);
        return fail(c, SPK_E_NOMEM, "column filter: device allocation failed");
    }
    r = cuda_check(c, cudaMemcpyAsync(d_idx, idx.data(), (size_t)Lk * 4, cudaMemcpyHostToDevice, c->stream), "H2D kept columns");
    if (!r) r = gather_columns_device(c, u.d_raw, n, L, d_idx, Lk, d_f);
    if (!r) r = cuda_check(c, cudaStreamSynchronize(c->stream), "sync");
    if (!r) r = set_alignment_impl(c, d_f, n, Lk, kpos.data(), cudaMemcpyDeviceToDevice);
    cudaFree(d_idx); cudaFree(d_f);
    return r;
}

int spk_neighbour_counts_part(spk_ctx* c, uint64_t min_identical, int quirk, uint32_t part, uint32_t nparts, uint64_t* counts) {
    if (!c || !counts) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "neighbour_counts: no alignment set");
    if (nparts == 0) nparts = 1;
    if (part >= nparts) return fail(c, SPK_E_INVALID, "neighbour_counts: part_index >= n_parts");
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = hamming_neighbour_counts(c, min_identical, quirk, part, nparts); if (r) return r;
    SPK_CUDA(c, cudaMemcpyAsync(counts, c->d_counts, (size_t)c->n * 8, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.hamming_ms, c->ev[0], c->ev[1]));
    return SPK_OK;
}

int spk_neighbour_counts(spk_ctx* c, uint64_t min_identical, int quirk, uint64_t* counts) {
    return spk_neighbour_counts_part(c, min_identical, quirk, 0, 1, counts);
}

int spk_set_weights(spk_ctx* c, const double* w) {
    if (!c || !w) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "set_weights: no alignment set");
    SPK_CUDA(c, cudaSetDevice(c->device));
    std::vector<double> hw(c->n_pad, 0.0);
    double tot = 0.0;
    for (uint32_t s = 0; s < c->n; ++s) { hw[s] = w[s]; tot += w[s]; }
    SPK_CUDA(c, cudaMemcpyAsync(c->d_w, hw.data(), (size_t)c->n_pad * 8, cudaMemcpyHostToDevice, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    c->w_total = tot;
    c->h_w.assign(w, w + c->n);
    c->i8_weights_valid = false; c->g_state = 0;
    c->operands_terms = 0; c->stats_valid = false;
    return SPK_OK;
}

int spk_mi_pairs(spk_ctx* c, const uint32_t* pi, const uint32_t* pj, uint64_t npairs, double pc, double* mi) {
    if (!c || (npairs && (!pi || !pj || !mi))) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "mi_pairs: no alignment set");
    if (npairs == 0) return SPK_OK;
    for (uint64_t k = 0; k < npairs; ++k) if (pi[k] >= c->L || pj[k] >= c->L) return fail(c, SPK_E_INVALID, "mi_pairs: column index out of range");
    SPK_CUDA(c, cudaSetDevice(c->device));
    const size_t ib = (npairs * 4 + 255) & ~(size_t)255, ob = npairs * 8;
    uint8_t* base = (uint8_t*)scratch(c, 2 * ib + ob);
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed");
    uint32_t* d_pi = (uint32_t*)base; uint32_t* d_pj = (uint32_t*)(base + ib); double* d_mi = (double*)(base + 2 * ib);
    SPK_CUDA(c, cudaMemcpyAsync(d_pi, pi, npairs * 4, cudaMemcpyHostToDevice, c->stream));
    SPK_CUDA(c, cudaMemcpyAsync(d_pj, pj, npairs * 4, cudaMemcpyHostToDevice, c->stream));
    int r = mi_pairs_exact(c, d_pi, d_pj, npairs, pc, d_mi); if (r) return r;
    SPK_CUDA(c, cudaMemcpyAsync(mi, d_mi, ob, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}

// host logic of determine_MI_threshold (include/SpydrPick.hpp:258-343)
int spk_determine_mi_threshold(spk_ctx* c, uint64_t n_values, uint64_t threshold_pairs, uint32_t iterations,
                               double pc, uint64_t seed, double* threshold) {
    if (!c || !threshold || iterations == 0) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "determine_mi_threshold: no alignment set");
    const uint64_t L = c->L;
    const uint64_t possible = L * (L - 1) / 2;
    const double pct = 1.0 - double(n_values) / double(possible);
    uint64_t tp = threshold_pairs;
    if (tp == 0) {                                           // SpydrPick.hpp:263-273
        tp = 100000;
        while (double(tp) - pct * double(tp) < 100.0 && tp < 500000) tp += 10000;
    }
    if (possible / 10 < tp) tp = possible / 10;              // SpydrPick.hpp:276-279
    if (tp == 0) return fail(c, SPK_E_INVALID, "alignment too small for threshold estimation");
    const uint64_t idx = (uint64_t)(pct * double(tp));
    std::vector<uint32_t> pi(tp), pj(tp);
    std::vector<double> mi(tp), thresholds;
    for (uint32_t it = 0; it < iterations; ++it) {
        std::mt19937 gen((uint32_t)(seed + it));
        std::uniform_int_distribution<uint64_t> dist(0, L - 1);
        std::unordered_set<uint64_t> seen; seen.reserve(tp * 2);
        uint64_t k = 0;
        while (k < tp) {
            uint64_t a = dist(gen), b = dist(gen);
            while (a == b) { a = dist(gen); b = dist(gen); }
            if (a > b) std::swap(a, b);
            if (seen.insert((a << 32) | b).second) { pi[k] = (uint32_t)a; pj[k] = (uint32_t)b; ++k; }
        }
        int r = spk_mi_pairs(c, pi.data(), pj.data(), tp, pc, mi.data()); if (r) return r;
        std::nth_element(mi.begin(), mi.begin() + idx, mi.end());
        thresholds.push_back(mi[idx]);
    }
    const size_t med = thresholds.size() / 2 - (thresholds.size() % 2 ? 0 : 1);
    std::nth_element(thresholds.begin(), thresholds.begin() + med, thresholds.end());
    *threshold = thresholds[med];
    return SPK_OK;
}

int spk_outlier_thresholds(const double* colmax, uint64_t L, double* outlier, double* extreme) {
    if (!colmax || !outlier || !extreme || L == 0) return SPK_E_INVALID;
    std::vector<double> v(colmax, colmax + L);
    std::sort(v.begin(), v.end());
    const double Q1 = v[L / 4 * 1], Q3 = v[L / 4 * 3];       // mi.hpp:280 (integer division first)
    *outlier = Q3 + 1.5 * (Q3 - Q1);
    *extreme = Q3 + 3.0 * (Q3 - Q1);
    return SPK_OK;
}

static int ensure_scan_buffers(Ctx* c, uint64_t max_edges, bool exhaustive) {
    uint64_t want_edges = max_edges;
    if (want_edges == 0) want_edges = 2 * std::min<uint64_t>(10000000ull, 100ull * c->L) + 65536;
    const uint64_t possible = (uint64_t)c->L * (c->L - 1) / 2;
    want_edges = std::min<uint64_t>(want_edges, possible + 1);
    uint64_t want_cand = want_edges + want_edges / 2 + (uint64_t)c->L * 64 + (1u << 20);
    want_cand = std::min<uint64_t>(want_cand, possible + 1);
    want_cand += (uint64_t)c->sm_count * 16 * 2 * 128;        // int8 engine: two partly used emission chunks per epilogue warp
    if (exhaustive) want_cand = std::max<uint64_t>(std::min<uint64_t>(possible + 1, 1ull << 24), (uint64_t)c->L);
    int r;
    if (c->edges_cap < want_edges) {
        if ((r = dev_alloc(c, c->d_edges, want_edges * sizeof(spk_edge), "edges"))) return r;
        if ((r = dev_alloc(c, c->d_wo, want_edges * sizeof(spk_edge), "edges_wo_gaps"))) return r;
        c->edges_cap = c->wo_cap = want_edges;
    }
    if (c->cand_cap < want_cand) {
        if ((r = dev_alloc(c, c->d_cand, want_cand * sizeof(Candidate), "candidates"))) return r;
        if ((r = dev_alloc(c, c->d_refined, want_cand * sizeof(spk_edge), "refined"))) return r;
        if ((r = dev_alloc(c, c->d_refined_wo, want_cand * 8, "refined_wo"))) return r;
        c->cand_cap = c->refined_cap = want_cand;
    }
    return SPK_OK;
}

int spk_mi_scan(spk_ctx* c, const spk_scan_params* pp, spk_scan_result* res) {
    if (!c || !pp || !res) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "mi_scan: no alignment set");
    spk_scan_params p = *pp;
    memset(res, 0, sizeof *res);
    SPK_CUDA(c, cudaSetDevice(c->device));
    if (!(p.pseudocount > 0.0)) return fail(c, SPK_E_INVALID, "mi_scan: pseudocount must be > 0 (0 makes the reference produce NaN via 0*log 0)");
    const int T = p.n_terms ? p.n_terms : 2;
    if (T < 1 || T > 3) return fail(c, SPK_E_INVALID, "mi_scan: n_terms must be 1..3");
    if (p.engine < 0 || p.engine > 4) return fail(c, SPK_E_INVALID, "mi_scan: engine must be 0..4");
    const bool use_i8 = (p.engine == 0 || p.engine == 3 || p.engine == 4);
    const int i8_wr = p.engine == 3 ? 2 : (p.engine == 4 ? 1 : i8_default_wr());
    // default band: measured filter errors are ~3e-6 (int8 engine 3), ~1e-5 (engine 4), 4e-5 (bf16 T=2), 4e-3 (T=1); when the band is left to
    // the library and a run's measured error exceeds band/2 the scan is repeated with a wider band (never silently accepted).
    double band = p.band > 0.0 ? p.band : (use_i8 ? (i8_wr == 2 ? 4e-5 : 2e-4) : (T == 1 ? 1e-2 : 2e-4));
    const bool band_auto = !(p.band > 0.0);
    const uint64_t circle = p.circle_size ? p.circle_size : c->L;
    if (p.n_parts == 0) p.n_parts = 1;
    if (p.part_index >= p.n_parts) return fail(c, SPK_E_INVALID, "mi_scan: part_index >= n_parts");
    if (c->dl_pending) { SPK_CUDA(c, cudaEventSynchronize(c->ev_dl)); c->dl_pending = false; }
    int r = ensure_scan_buffers(c, p.max_edges, p.engine == 1); if (r) return r;

    unsigned long long h_counters[8] = {0};
    uint64_t ne = 0, nw = 0;
    for (int attempt = 0;; ++attempt) {
    SPK_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 8 * sizeof(unsigned long long), c->stream));
    fill_u64_kernel<<<(c->L_pad + 255) / 256, 256, 0, c->stream>>>((unsigned long long*)c->d_colmax, f64_ordered(-DBL_MAX), c->L_pad);
    c->tm.kernel_launches += 1;
    c->tm.gemm_ms = c->tm.refine_ms = c->tm.compact_ms = c->tm.sort_ms = 0.f; c->tm.gemm_tiles = 0;

    if (p.engine == 1) {
        r = mi_scan_exact(c, p, circle); if (r) return r;
    } else {
        if (use_i8) {
            // the refine's plane build only needs the weights: queued first, it runs on the GPU while the host prepares the filter
            if (attempt == 0) { r = mi_refine_prepare(c, p.pseudocount); if (r) return r; }
            r = mi_i8_filter(c, p, band, circle, i8_wr); if (r) return r;
        } else {
            r = prep_column_stats(c, p.pseudocount); if (r) return r;
            r = prep_operands(c, T); if (r) return r;
            r = mi_gemm_filter(c, p, T, band, circle); if (r) return r;
        }
        SPK_CUDA(c, cudaMemcpyAsync(h_counters, c->d_counters, sizeof h_counters, cudaMemcpyDeviceToHost, c->stream));
        SPK_CUDA(c, cudaStreamSynchronize(c->stream));
        if (h_counters[4]) return fail(c, SPK_E_OVERFLOW, "candidate buffer overflow (%llu candidates > capacity %llu): raise max_edges or the threshold",
                                       h_counters[0], (unsigned long long)c->cand_cap);
        uint64_t ncand = h_counters[0];                  // int8 engine: reserved slots, a few of them invalid-marked holes
        SPK_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
        r = mi_refine_candidates(c, ncand, p.pseudocount, p.threshold, &ncand); if (r) return r;
        res->n_candidates = ncand;
        SPK_CUDA(c, cudaEventRecord(c->ev[5], c->stream));
        r = mi_compact(c, ncand, p, circle); if (r) return r;
        SPK_CUDA(c, cudaEventRecord(c->ev[6], c->stream));
        SPK_CUDA(c, cudaEventSynchronize(c->ev[6]));
        SPK_CUDA(c, cudaEventElapsedTime(&c->tm.refine_ms, c->ev[4], c->ev[5]));
        SPK_CUDA(c, cudaEventElapsedTime(&c->tm.compact_ms, c->ev[5], c->ev[6]));
    }
    ordered_to_double_kernel<<<(c->L_pad + 255) / 256, 256, 0, c->stream>>>((unsigned long long*)c->d_colmax, c->L_pad);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaMemcpyAsync(h_counters, c->d_counters, sizeof h_counters, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    if (h_counters[4]) return fail(c, SPK_E_OVERFLOW, "edge buffer overflow (%llu edges > capacity %llu): raise max_edges", h_counters[1],
                                   (unsigned long long)c->edges_cap);
    ne = h_counters[1]; nw = h_counters[2];
    res->max_filter_error = h_counters[3] ? f64_from_ordered(h_counters[3]) : 0.0;
    if (p.engine != 1 && res->max_filter_error > 0.5 * band) {
        if (band_auto && attempt < 3) {
            const double wider = std::max(4.0 * res->max_filter_error, 4.0 * band);
            if (c->verbose) fprintf(stderr, "spydrpick_b200: filter error %.3g > band/2 (band %.3g): rescanning with band %.3g\n",
                                    res->max_filter_error, band, wider);
            band = wider;
            continue;
        }
        return fail(c, SPK_E_PRECISION, "fp32 filter error %.3g exceeds half the safety band %.3g: raise band (or n_terms with engine 2)",
                    res->max_filter_error, band);
    }
    break;
    }
    // pinned mirrors: D2H at full PCIe rate, no page-fault pass over a fresh std::vector
    auto grow = [&](spk_edge*& buf, uint64_t& cap, uint64_t want) -> int {
        if (want <= cap) return SPK_OK;
        if (buf) { cudaFreeHost(buf); buf = nullptr; cap = 0; }
        const uint64_t newcap = want + want / 4 + 1024;
        if (cudaHostAlloc((void**)&buf, newcap * sizeof(spk_edge), cudaHostAllocDefault) != cudaSuccess) {
            cudaGetLastError(); return fail(c, SPK_E_NOMEM, "pinned host allocation failed (%llu edges)", (unsigned long long)newcap);
        }
        cap = newcap; return SPK_OK;
    };
    // colmax is final after compaction: read back BEFORE the bulk downloads are queued (one D2H copy engine, in issue order)
    c->h_colmax.resize(c->L);
    SPK_CUDA(c, cudaMemcpyAsync(c->h_colmax.data(), c->d_colmax, (size_t)c->L * 8, cudaMemcpyDeviceToHost, c->stream));
    bool wo_in_flight = false;
    if (p.keep_on_device == 2 && nw) {
        // the MI-without-gaps list is final after compaction (the sort only touches the edge list): its download
        // starts now on the side stream and runs under the edge sort
        if ((r = grow(c->hp_wo, c->hp_wo_cap, nw))) return r;
        SPK_CUDA(c, cudaEventRecord(c->ev_ready, c->stream));
        SPK_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_ready, 0));
        SPK_CUDA(c, cudaMemcpyAsync(c->hp_wo, c->d_wo, nw * sizeof(spk_edge), cudaMemcpyDeviceToHost, c->copy_stream));
        wo_in_flight = true;
    }
    c->d_edges_sorted_n = 0;
    if (p.sort_desc) { r = sort_edges_device(c, c->d_edges, ne); if (r) return r; c->d_edges_sorted_n = ne; }
    res->n_edges = ne; res->n_wo_gaps = nw;
    res->d_edges = c->d_edges; res->d_edges_wo_gaps = c->d_wo; res->d_colmax = c->d_colmax;
    c->host_edges_n = 0;
    if (p.keep_on_device != 1) {
        if ((r = grow(c->hp_edges, c->hp_edges_cap, ne))) return r;
        if ((r = grow(c->hp_wo, c->hp_wo_cap, nw))) return r;
        // keep_on_device == 2: the copies run on a side stream and overlap whatever the caller launches next (ARACNE on the
        // device copy); the host pointers are returned now, their contents are valid after spk_wait_download().
        cudaStream_t cs = c->stream;
        if (p.keep_on_device == 2) {
            SPK_CUDA(c, cudaEventRecord(c->ev_ready, c->stream));
            SPK_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_ready, 0));
            cs = c->copy_stream;
        }
        if (ne) SPK_CUDA(c, cudaMemcpyAsync(c->hp_edges, c->d_edges, ne * sizeof(spk_edge), cudaMemcpyDeviceToHost, cs));
        if (nw && !wo_in_flight) SPK_CUDA(c, cudaMemcpyAsync(c->hp_wo, c->d_wo, nw * sizeof(spk_edge), cudaMemcpyDeviceToHost, cs));
        if (p.keep_on_device == 2) { SPK_CUDA(c, cudaEventRecord(c->ev_dl, cs)); c->dl_pending = true; }
        if (p.sort_desc) c->host_edges_n = ne;
    }
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    res->colmax = c->h_colmax.data();
    if (p.keep_on_device != 1) { res->edges = c->hp_edges; res->edges_wo_gaps = c->hp_wo; }
    return SPK_OK;
}

int spk_sort_edges_device(spk_ctx* c, void* d_edges, uint64_t n) {
    if (!c || (!d_edges && n)) return SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = sort_edges_device(c, (spk_edge*)d_edges, n); if (r) return r;
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}

int spk_sort_edges(spk_ctx* c, spk_edge* edges, uint64_t n) {
    if (!c || (!edges && n)) return SPK_E_INVALID;
    if (n < 2) return SPK_OK;
    SPK_CUDA(c, cudaSetDevice(c->device));
    spk_edge* d = nullptr;
    if (cudaMalloc(&d, n * sizeof(spk_edge)) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "sort: device allocation failed"); }
    int r = cuda_check(c, cudaMemcpyAsync(d, edges, n * sizeof(spk_edge), cudaMemcpyHostToDevice, c->stream), "H2D edges");
    if (!r) r = sort_edges_device(c, d, n);
    if (!r) r = cuda_check(c, cudaMemcpyAsync(edges, d, n * sizeof(spk_edge), cudaMemcpyDeviceToHost, c->stream), "D2H edges");
    if (!r) r = cuda_check(c, cudaStreamSynchronize(c->stream), "sync");
    cudaFree(d);
    return r;
}

int spk_merge_sorted_runs_device(spk_ctx* c, void* d_edges, const uint64_t* run_counts, uint32_t n_runs) {
    if (!c || !run_counts) return SPK_E_INVALID;
    uint64_t n = 0;
    for (uint32_t r = 0; r < n_runs; ++r) n += run_counts[r];
    if (n && !d_edges) return SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    return merge_sorted_runs_device(c, (spk_edge*)d_edges, run_counts, n_runs);
}

int spk_aracne_device(spk_ctx* c, const void* d_edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, void* d_flags) {
    if (!c || (n && (!d_edges || !d_flags))) return SPK_E_INVALID;
    if (nparts == 0) nparts = 1;
    if (part >= nparts) return fail(c, SPK_E_INVALID, "aracne: part_index >= n_parts");
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = aracne_device(c, (const spk_edge*)d_edges, n, eps, part, nparts, (uint8_t*)d_flags); if (r) return r;
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}

int spk_aracne(spk_ctx* c, const spk_edge* edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, uint8_t* flags) {
    if (!c || (n && (!edges || !flags))) return SPK_E_INVALID;
    if (n == 0) return SPK_OK;
    if (nparts == 0) nparts = 1;
    if (part >= nparts) return fail(c, SPK_E_INVALID, "aracne: part_index >= n_parts");
    SPK_CUDA(c, cudaSetDevice(c->device));
    if (edges == c->hp_edges && n == c->host_edges_n && c->d_edges) {
        // the caller hands back the sorted edge list of the last scan unchanged (the reference's
        // run_ARACNE(network) right after get_MI_network): the device copy is still valid, skip the upload.
        if (c->d_flags_cap < n) {                                // (aracne_device owns the context scratch buffer)
            if (c->d_flags_buf) { cudaFree(c->d_flags_buf); c->d_flags_buf = nullptr; c->d_flags_cap = 0; }
            const uint64_t cap = n + n / 4 + 4096;
            if (cudaMalloc(&c->d_flags_buf, cap) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "aracne: device allocation failed"); }
            c->d_flags_cap = cap;
        }
        uint8_t* f0 = c->d_flags_buf;
        // whole list evaluated: every flag is written; a partition leaves the others as the caller passed them
        int r0 = (nparts <= 1) ? cuda_check(c, cudaMemsetAsync(f0, 0, n, c->stream), "clear flags")
                               : cuda_check(c, cudaMemcpyAsync(f0, flags, n, cudaMemcpyHostToDevice, c->stream), "H2D flags");
        if (!r0) r0 = aracne_device(c, c->d_edges, n, eps, part, nparts, f0);
        if (!r0) r0 = cuda_check(c, cudaMemcpyAsync(flags, f0, n, cudaMemcpyDeviceToHost, c->stream), "D2H flags");
        if (!r0) r0 = cuda_check(c, cudaStreamSynchronize(c->stream), "sync");
        return r0;
    }
    spk_edge* d = nullptr; uint8_t* f = nullptr;
    if (cudaMalloc(&d, n * sizeof(spk_edge)) != cudaSuccess || cudaMalloc(&f, n) != cudaSuccess) {
        cudaGetLastError(); if (d) cudaFree(d);
        return fail(c, SPK_E_NOMEM, "aracne: device allocation failed");
    }
    int r = cuda_check(c, cudaMemcpyAsync(d, edges, n * sizeof(spk_edge), cudaMemcpyHostToDevice, c->stream), "H2D edges");
    if (!r) r = cuda_check(c, cudaMemcpyAsync(f, flags, n, cudaMemcpyHostToDevice, c->stream), "H2D flags");
    if (!r) r = aracne_device(c, d, n, eps, part, nparts, f);
    if (!r) r = cuda_check(c, cudaMemcpyAsync(flags, f, n, cudaMemcpyDeviceToHost, c->stream), "D2H flags");
    if (!r) r = cuda_check(c, cudaStreamSynchronize(c->stream), "sync");
    cudaFree(d); cudaFree(f);
    return r;
}

int spk_wait_download(spk_ctx* c) {
    if (!c) return SPK_E_INVALID;
    if (c->dl_pending) { SPK_CUDA(c, cudaEventSynchronize(c->ev_dl)); c->dl_pending = false; }
    return SPK_OK;
}

int spk_get_timings(spk_ctx* c, spk_timings* out) {
    if (!c || !out) return SPK_E_INVALID;
    *out = c->tm;
    return SPK_OK;
}

int spk_debug_tile_counts(spk_ctx* c, uint32_t ib, uint32_t jb, int32_t n_terms, float* counts) {
    if (!c || !counts) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "debug_tile_counts: no alignment set");
    const int T = n_terms ? n_terms : 2;
    if (T < 1 || T > 3) return fail(c, SPK_E_INVALID, "n_terms must be 1..3");
    if (ib >= (c->L + TILE_I - 1) / TILE_I || jb >= c->L_pad / TILE_J) return fail(c, SPK_E_INVALID, "tile out of range");
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = prep_column_stats(c, 0.5); if (r) return r;
    r = prep_operands(c, T); if (r) return r;
    const size_t bytes = (size_t)TILE_I * TILE_J * 16 * sizeof(float);
    float* d_out = nullptr;
    if (cudaMalloc(&d_out, bytes) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "debug buffer"); }
    cudaMemsetAsync(d_out, 0xff, bytes, c->stream);
    r = mi_gemm_debug_tile(c, ib, jb, T, d_out);
    if (!r) r = cuda_check(c, cudaMemcpy(counts, d_out, bytes, cudaMemcpyDeviceToHost), "D2H debug tile");
    cudaFree(d_out);
    return r;
}

int spk_debug_i8_mode(void) { return i8_default_wr(); }

int spk_debug_i8_layout(spk_ctx* c, uint32_t* ntiles3, uint32_t* col_start3, uint32_t* perm, uint32_t* info, uint64_t perm_cap,
                        uint32_t* q, double* delta) {
    if (!c || !ntiles3 || !col_start3) return SPK_E_INVALID;
    if (!c->d_states || c->i8.NT == 0) return fail(c, SPK_E_INVALID, "debug_i8_layout: no alignment set");
    SPK_CUDA(c, cudaSetDevice(c->device));
    for (int g = 0; g < 3; ++g) { ntiles3[g] = c->i8.ntiles[g]; col_start3[g] = c->i8.col_start[g]; }
    if (perm || info) {
        if (perm_cap < c->i8.Lperm) return fail(c, SPK_E_INVALID, "debug_i8_layout: perm_cap < %u", c->i8.Lperm);
        if (perm) SPK_CUDA(c, cudaMemcpy(perm, c->d_i8_perm, (size_t)c->i8.Lperm * 4, cudaMemcpyDeviceToHost));
        if (info) SPK_CUDA(c, cudaMemcpy(info, c->d_i8_info, (size_t)c->i8.Lperm * 4, cudaMemcpyDeviceToHost));
    }
    if (q) {
        if (!c->i8_weights_valid) return fail(c, SPK_E_INVALID, "debug_i8_layout: weights not quantised yet (run a scan or debug tile first)");
        SPK_CUDA(c, cudaMemcpy(q, c->d_q, (size_t)c->n * 4, cudaMemcpyDeviceToHost));
    }
    if (delta) *delta = c->i8_delta;
    return SPK_OK;
}

int spk_debug_i8_tile(spk_ctx* c, uint32_t ti, uint32_t tj, uint32_t* tables /* [16384] */) {
    if (!c || !tables) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "debug_i8_tile: no alignment set");
    SPK_CUDA(c, cudaSetDevice(c->device));
    uint32_t* d_out = nullptr;
    const size_t bytes = 16384 * sizeof(uint32_t);
    if (cudaMalloc(&d_out, bytes) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "debug buffer"); }
    cudaMemsetAsync(d_out, 0xff, bytes, c->stream);
    int r = mi_i8_debug_tile(c, ti, tj, 0.5, d_out);
    if (!r) r = cuda_check(c, cudaMemcpy(tables, d_out, bytes, cudaMemcpyDeviceToHost), "D2H debug tile");
    cudaFree(d_out);
    return r;
}

int spk_host_alloc(size_t bytes, void** out) {
    if (!out) return SPK_E_INVALID;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes ? bytes : 16, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return SPK_E_NOMEM; }
    return SPK_OK;
}
void spk_host_free(void* p) { if (p) cudaFreeHost(p); }

} // extern "C"
