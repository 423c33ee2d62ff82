// spk_cabi.cu -- the C-ABI (include/spydrpick_b200.h): context, buffers, stage orchestration.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdarg>
#include <cstring>
#include <random>
#include <thread>
#include <unordered_set>
#include <cub/cub.cuh>
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
    free_dev(c->d_merge); free_dev(c->d_wo_all); free_dev(c->d_audit); free_dev(c->d_xcnt);
    c->cand_cap = c->edges_cap = c->wo_cap = c->refined_cap = c->merge_cap = c->wo_all_cap = c->audit_cap = 0;
    c->d_edges_view = nullptr;
}

// turns the outstanding (start, stop) event pairs into spk_timings fields; the stream must be idle
void resolve_timings(Ctx* c) {
    struct { uint32_t bit; int a, b; float* out; } tab[] = {
        {TMP_UPLOAD, EV_UP0, EV_UP1, &c->tm.upload_ms}, {TMP_PREP, EV_UP1, EV_PREP1, &c->tm.prep_ms}, {TMP_K1, EV_K1_0, EV_K1_1, &c->tm.hamming_ms},
        {TMP_EXPAND, EV_EXP0, EV_EXP1, &c->tm.expand_ms}, {TMP_STATS, EV_EXP1, EV_STAT1, &c->tm.stats_ms}, {TMP_GEMM, EV_GEMM0, EV_GEMM1, &c->tm.gemm_ms},
        {TMP_SEED, EV_GEMM0, EV_SEED1, &c->tm.seed_ms}, {TMP_REFINE, EV_REF0, EV_REF1, &c->tm.refine_ms}, {TMP_COMPACT, EV_REF1, EV_CMP1, &c->tm.compact_ms},
        {TMP_SORT, EV_SORT0, EV_SORT1, &c->tm.sort_ms}, {TMP_XCH, EV_XCH0, EV_XCH1, &c->tm.exchange_ms}, {TMP_ARB, EV_AR0, EV_AR1, &c->tm.aracne_build_ms},
        {TMP_AR, EV_AR1, EV_AR2, &c->tm.aracne_ms}, {TMP_PAIRS, EV_PAIRS0, EV_PAIRS1, &c->tm.pairs_ms}, {TMP_AUDIT, EV_AUD0, EV_AUD1, &c->tm.audit_ms},
        {TMP_THR, EV_THR0, EV_THR1, &c->tm.threshold_ms}};
    for (auto& t : tab) {
        if (!(c->tm_pending & t.bit)) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, c->ev[t.a], c->ev[t.b]) == cudaSuccess) *t.out = ms; else cudaGetLastError();
    }
    c->tm_pending = 0;
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
// w[s] = 1 / count[s] (multiplicity is always 1: Alignment_StateVector_weights.hpp:349); pad samples 0
__global__ void weights_from_counts_kernel(const unsigned long long* __restrict__ counts, double* __restrict__ w, uint32_t n, uint32_t n_pad) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_pad) w[i] = i < n ? 1.0 / (double)counts[i] : 0.0;
}

} // namespace spk

using namespace spk;

extern "C" {

const char* spk_version(void) { return "spydrpick_b200 0.2 (sm_100a)"; }

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
    c->d_edges_view = nullptr;
    cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
    cudaEventCreateWithFlags(&c->ev_ready, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->ev_dl, cudaEventDisableTiming);
    if (cudaMalloc(&c->d_counters, 16 * sizeof(unsigned long long)) != cudaSuccess || cudaMalloc(&c->d_tilemap, 1 << 20) != cudaSuccess) {
        delete c; return fail(nullptr, SPK_E_NOMEM, "context allocation failed");
    }
    *out = c;
    return SPK_OK;
}

void spk_destroy(spk_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    comm_free(c);
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
    // multi-GPU, host source: every rank uploads its block of rows over its own PCIe link, the blocks are all-gathered over
    // NVLink ("alignment broadcast once") -- the row buffer is padded to world * rows_per_rank rows for that
    const uint32_t per = (c->world > 1) ? (n + (uint32_t)c->world - 1) / (uint32_t)c->world : n;
    if (!same_shape) {
        free_alignment(c);
        c->n = n; c->L = L;
        c->n_pad = (n + 127) / 128 * 128;                    // K blocks: 64 samples (bf16 engine), 128 (int8 engine)
        c->L_pad = (L + TILE_J - 1) / TILE_J * TILE_J;
        c->W32 = (L + 31) / 32;
        int r;
        if ((r = dev_alloc(c, c->d_states, (size_t)per * (size_t)(c->world > 1 ? c->world : 1) * L, "states"))) return r;
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
    c->operands_terms = 0; c->stats_valid = false; c->host_edges_n = 0; c->g_state = 0; c->d_edges_sorted_n = 0; c->i8_seed_valid = false;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_UP0], c->stream));
    if (c->world > 1 && kind == cudaMemcpyHostToDevice) {
        const uint32_t r0 = std::min(n, per * (uint32_t)c->rank), r1 = std::min(n, r0 + per);
        if (r1 > r0) SPK_CUDA(c, cudaMemcpyAsync(c->d_states + (size_t)r0 * L, states + (size_t)r0 * L, (size_t)(r1 - r0) * L, kind, c->stream));
        int rx = comm_allgather_inplace(c, c->d_states, (size_t)per * L); if (rx) return rx;
    } else {
        SPK_CUDA(c, cudaMemcpyAsync(c->d_states, states, (size_t)n * L, kind, c->stream));
    }
    std::vector<uint64_t> hp(c->L_pad, 0);
    for (uint32_t k = 0; k < L; ++k) hp[k] = pos ? pos[k] : k;
    SPK_CUDA(c, cudaMemcpyAsync(c->d_pos, hp.data(), (size_t)c->L_pad * 8, cudaMemcpyHostToDevice, c->stream));
    std::vector<double> ones(c->n_pad, 0.0);
    for (uint32_t s = 0; s < n; ++s) ones[s] = 1.0;
    SPK_CUDA(c, cudaMemcpyAsync(c->d_w, ones.data(), (size_t)c->n_pad * 8, cudaMemcpyHostToDevice, c->stream));
    c->w_total = (double)n;
    c->h_w.assign(n, 1.0);
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_UP1], c->stream));
    int r = prep_alignment(c); if (r) return r;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_PREP1], c->stream));
    c->tm_pending |= TMP_UPLOAD | TMP_PREP;
    r = i8_build_layout(c); if (r) return r;             // reads the presence masks back: the stream is idle afterwards
    resolve_timings(c);
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

int spk_neighbour_counts_part(spk_ctx* c, uint64_t min_identical, int quirks, uint32_t part, uint32_t nparts, uint64_t* counts) {
    if (!c || !counts) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "neighbour_counts: no alignment set");
    if (nparts == 0) nparts = 1;
    if (part >= nparts) return fail(c, SPK_E_INVALID, "neighbour_counts: part_index >= n_parts");
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = hamming_neighbour_counts(c, min_identical, quirks, part, nparts); if (r) return r;
    SPK_CUDA(c, cudaMemcpyAsync(counts, c->d_counts, (size_t)c->n * 8, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    resolve_timings(c);
    return SPK_OK;
}

int spk_neighbour_counts(spk_ctx* c, uint64_t min_identical, int quirks, uint64_t* counts) {
    if (c && c->world > 1) {
        // collective: this rank's share of the sample-tile pairs, partial counts summed over NVLink
        if (!counts) return SPK_E_INVALID;
        if (!c->d_states) return fail(c, SPK_E_INVALID, "neighbour_counts: no alignment set");
        SPK_CUDA(c, cudaSetDevice(c->device));
        int r = hamming_neighbour_counts(c, min_identical, quirks, (uint32_t)c->rank, (uint32_t)c->world); if (r) return r;
        r = comm_allreduce_sum_u64(c, (unsigned long long*)c->d_counts, c->n); if (r) return r;
        SPK_CUDA(c, cudaMemcpyAsync(counts, c->d_counts, (size_t)c->n * 8, cudaMemcpyDeviceToHost, c->stream));
        SPK_CUDA(c, cudaStreamSynchronize(c->stream));
        resolve_timings(c);
        return SPK_OK;
    }
    return spk_neighbour_counts_part(c, min_identical, quirks, 0, 1, counts);
}

static int set_weights_host(spk_ctx* c, const double* w) {
    std::vector<double> hw(c->n_pad, 0.0);
    double tot = 0.0;
    for (uint32_t s = 0; s < c->n; ++s) { hw[s] = w[s]; tot += w[s]; }
    SPK_CUDA(c, cudaMemcpyAsync(c->d_w, hw.data(), (size_t)c->n_pad * 8, cudaMemcpyHostToDevice, c->stream));   // pageable source: staged before the call returns
    c->w_total = tot;
    c->h_w.assign(w, w + c->n);
    c->i8_weights_valid = false; c->g_state = 0; c->i8_seed_valid = false;
    c->operands_terms = 0; c->stats_valid = false;
    return SPK_OK;
}

int spk_set_weights(spk_ctx* c, const double* w) {
    if (!c || !w) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "set_weights: no alignment set");
    SPK_CUDA(c, cudaSetDevice(c->device));
    return set_weights_host(c, w);
}

// K1 + cache_sample_weights' computed branch in one call: counts -> w = 1/count -> optional rescale -> context weights.
int spk_reweight(spk_ctx* c, uint64_t min_identical, int quirks, int rescale, double* weights_out, uint64_t* counts_out) {
    if (!c) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "reweight: no alignment set");
    SPK_CUDA(c, cudaSetDevice(c->device));
    const uint32_t part = c->world > 1 ? (uint32_t)c->rank : 0u, nparts = c->world > 1 ? (uint32_t)c->world : 1u;
    int r = hamming_neighbour_counts(c, min_identical, quirks, part, nparts); if (r) return r;
    r = comm_allreduce_sum_u64(c, (unsigned long long*)c->d_counts, c->n); if (r) return r;
    std::vector<uint64_t> hc(c->n);
    SPK_CUDA(c, cudaMemcpyAsync(hc.data(), c->d_counts, (size_t)c->n * 8, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    resolve_timings(c);
    std::vector<double> w(c->n);
    for (uint32_t s = 0; s < c->n; ++s) w[s] = 1.0 / double(hc[s]);                      // weights.hpp:349
    if (rescale) {                                                                     // (w * n) / n_eff, Alignment_utility.hpp:233,245
        double n_eff = 0.0; for (double x : w) n_eff += x;
        for (double& x : w) x = x * double(c->n) / n_eff;
    }
    if (counts_out) std::memcpy(counts_out, hc.data(), (size_t)c->n * 8);
    if (weights_out) std::memcpy(weights_out, w.data(), (size_t)c->n * 8);
    return set_weights_host(c, w.data());
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
    resolve_timings(c);
    return SPK_OK;
}

} // extern "C"

namespace spk {
// ---- threshold estimator, device side -----------------------------------------------------------------------------
// SPK_SAMPLER_DEVICE: `count` DISTINCT pairs per iteration without any rejection loop -- slot k of iteration `it` is pair
// number perm(k) of the L(L-1)/2 upper-triangle pairs, perm = a keyed 4-round Feistel permutation of [0, 4^h) walked until
// it lands inside [0, P) (cycle walking keeps it a bijection on [0, P)). Same distribution as the reference's sample_pairs()
// (a uniformly random set of distinct pairs, include/SpydrPick.hpp:171-207), different random numbers -- the reference's are
// clock-seeded anyway.
__device__ __forceinline__ uint32_t feistel_round(uint32_t x, uint64_t key, uint32_t round) {
    uint64_t z = (uint64_t)x * 0x9E3779B97F4A7C15ull + key + (uint64_t)round * 0xBF58476D1CE4E5B9ull;
    z ^= z >> 31; z *= 0x94D049BB133111EBull; z ^= z >> 29; z *= 0xD6E8FEB86659FD93ull; z ^= z >> 32;
    return (uint32_t)z;
}
__global__ void sample_pairs_kernel(uint32_t* __restrict__ pi, uint32_t* __restrict__ pj, uint64_t count, uint32_t iterations,
                                    uint32_t it0, uint32_t it_stride, uint64_t L, uint64_t P, uint32_t half_bits, uint64_t seed) {
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count * iterations) return;
    const uint32_t slot_it = (uint32_t)(t / count);
    const uint64_t k = t % count;
    const uint64_t key = seed + 0x632BE59BD9B4E019ull * (uint64_t)(it0 + slot_it * it_stride + 1);
    const uint32_t hmask = (half_bits >= 32) ? 0xffffffffu : ((1u << half_bits) - 1u);
    uint64_t x = k;
    do {
        uint32_t l = (uint32_t)(x >> half_bits) & hmask, r = (uint32_t)x & hmask;
#pragma unroll
        for (uint32_t rd = 0; rd < 4; ++rd) { const uint32_t f = feistel_round(r, key, rd) & hmask; const uint32_t nl = r; r = l ^ f; l = nl; }
        x = ((uint64_t)l << half_bits) | r;
    } while (x >= P);
    // pair number x -> (i, j), i < j: row i holds L-1-i pairs
    const double b = 2.0 * (double)L - 1.0;
    uint64_t i = (uint64_t)((b - sqrt(fmax(b * b - 8.0 * (double)x, 0.0))) * 0.5);
    if (i > L - 2) i = L - 2;
    auto row_start = [&](uint64_t q) { return q * (2 * L - q - 1) / 2; };
    while (i > 0 && row_start(i) > x) --i;
    while (row_start(i + 1) <= x) ++i;
    pi[t] = (uint32_t)i; pj[t] = (uint32_t)(i + 1 + (x - row_start(i)));
}
__global__ void segment_ids_kernel(uint32_t* __restrict__ seg, unsigned long long total, unsigned long long per) {
    const unsigned long long p = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < total) seg[p] = (uint32_t)(p / per);
}
__global__ void pick_kth_kernel(const double* __restrict__ sorted, uint64_t count, uint64_t idx, uint32_t iterations, double* __restrict__ out) {
    const uint32_t it = blockIdx.x * blockDim.x + threadIdx.x;
    if (it < iterations) out[it] = sorted[(uint64_t)it * count + idx];
}
}  // namespace spk

extern "C" {

// determine_MI_threshold (include/SpydrPick.hpp:258-343): per iteration, the MI of `threshold_pairs` distinct random pairs and
// the element at index pct * threshold_pairs of their ascending order; the result is the lower median over the iterations.
// All iterations go through ONE K2b launch and one segmented sort on the device; only `iterations` doubles come back.
int spk_determine_mi_threshold_ex(spk_ctx* c, uint64_t n_values, uint64_t threshold_pairs, uint32_t iterations,
                                  double pc, uint64_t seed, int sampler, double* threshold) {
    if (!c || !threshold || iterations == 0) return SPK_E_INVALID;
    if (!c->d_states) return fail(c, SPK_E_INVALID, "determine_mi_threshold: no alignment set");
    if (sampler != SPK_SAMPLER_REFERENCE && sampler != SPK_SAMPLER_DEVICE) return fail(c, SPK_E_INVALID, "determine_mi_threshold: unknown sampler");
    const uint64_t L = c->L;
    const uint64_t possible = L * (L - 1) / 2;
    const double pct = 1.0 - double(n_values) / double(possible);
    if (!(pct > 0.0)) return fail(c, SPK_E_INVALID, "determine_mi_threshold: n_values (%llu) is not smaller than the number of pairs (%llu): every pair would be saved "
                                  "(the reference's estimator indexes out of range here); use threshold 0", (unsigned long long)n_values, (unsigned long long)possible);
    uint64_t tp = threshold_pairs;
    if (tp == 0) {                                           // SpydrPick.hpp:263-273
        tp = 100000;
        while (double(tp) - pct * double(tp) < 100.0 && tp < 500000) tp += 10000;
    }
    if (possible / 10 < tp) tp = possible / 10;              // SpydrPick.hpp:276-279
    if (tp == 0) return fail(c, SPK_E_INVALID, "alignment too small for threshold estimation");
    uint64_t idx = (uint64_t)(pct * double(tp));
    if (idx >= tp) idx = tp - 1;
    SPK_CUDA(c, cudaSetDevice(c->device));
    // multi-GPU: iterations are dealt to the ranks, the per-iteration thresholds are summed into place (zeros elsewhere)
    const uint32_t world = (uint32_t)std::max(1, c->world), rank = (uint32_t)c->rank;
    std::vector<uint32_t> my_its;
    for (uint32_t it = rank; it < iterations; it += world) my_its.push_back(it);
    const uint32_t nit = (uint32_t)my_its.size();
    int r = mi_refine_prepare(c, pc); if (r) return r;       // every rank (the plane build may be a collective)
    const uint64_t total = tp * nit;
    if (total > 0x7fffffffull) return fail(c, SPK_E_INVALID, "determine_mi_threshold: too many sampled pairs");
    const size_t ib = (total * 4 + 255) & ~(size_t)255, ob = (total * 8 + 255) & ~(size_t)255;
    size_t tmpb = 0;
    std::vector<int> h_off(nit + 1);
    for (uint32_t k = 0; k <= nit; ++k) h_off[k] = (int)(tp * k);
    // per-iteration ascending order from two GLOBAL stable radix sorts (by value, then by iteration): a segmented sort runs one
    // block per segment -- 1.1 ms for 10 x 100,000 values against 0.2 ms this way
    uint32_t seg_bits = 1; while ((1u << seg_bits) < std::max(nit, 2u)) ++seg_bits;
    if (nit) {
        size_t t1 = 0, t2 = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, t1, (const double*)nullptr, (double*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)total, 0, 64, c->stream);
        cub::DeviceRadixSort::SortPairs(nullptr, t2, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const double*)nullptr, (double*)nullptr, (int)total, 0, (int)seg_bits, c->stream);
        tmpb = std::max(t1, t2);
    }
    const size_t offb = ((size_t)(nit + 1) * 4 + 255) & ~(size_t)255, thrb = ((size_t)iterations * 8 + 255) & ~(size_t)255;
    const size_t workb = (mi_pairs_segmented_scratch(total, nit) + 255) & ~(size_t)255;
    uint8_t* base = (uint8_t*)scratch(c, 2 * ib + 2 * ob + offb + thrb + ((tmpb + 255) & ~(size_t)255) + 256 + workb + 2 * ib + ob);
    if (!base) return fail(c, SPK_E_NOMEM, "scratch allocation failed");
    uint32_t* d_pi = (uint32_t*)base; uint32_t* d_pj = (uint32_t*)(base + ib);
    double* d_mi = (double*)(base + 2 * ib); double* d_sorted = (double*)(base + 2 * ib + ob);
    int* d_off = (int*)(base + 2 * ib + 2 * ob); double* d_thr = (double*)(base + 2 * ib + 2 * ob + offb);
    void* d_tmp = base + 2 * ib + 2 * ob + offb + thrb;
    uint8_t* d_work = base + 2 * ib + 2 * ob + offb + thrb + ((tmpb + 255) & ~(size_t)255) + 256;
    uint32_t* d_seg = (uint32_t*)(d_work + workb); uint32_t* d_seg2 = (uint32_t*)(d_work + workb + ib);
    double* d_byval = (double*)(d_work + workb + 2 * ib);
    SPK_CUDA(c, cudaMemsetAsync(d_thr, 0, (size_t)iterations * 8, c->stream));
    if (nit) {
        if (sampler == SPK_SAMPLER_REFERENCE) {
            // the reference's sample_pairs() with an explicit seed (it seeds from the clock, SpydrPick.hpp:178): one host thread per iteration
            std::vector<uint32_t> pi(total), pj(total);
            std::vector<std::thread> th;
            for (uint32_t k = 0; k < nit; ++k)
                th.emplace_back([&, k]() {
                    std::mt19937 gen((uint32_t)(seed + my_its[k]));
                    std::uniform_int_distribution<uint64_t> dist(0, L - 1);
                    std::unordered_set<uint64_t> seen; seen.reserve(tp * 2);
                    uint64_t q = 0; uint32_t* oi = pi.data() + (size_t)k * tp; uint32_t* oj = pj.data() + (size_t)k * tp;
                    while (q < tp) {
                        uint64_t a = dist(gen), b = dist(gen);
                        while (a == b) { a = dist(gen); b = dist(gen); }
                        if (a > b) std::swap(a, b);
                        if (seen.insert((a << 32) | b).second) { oi[q] = (uint32_t)a; oj[q] = (uint32_t)b; ++q; }
                    }
                });
            for (auto& t : th) t.join();
            SPK_CUDA(c, cudaMemcpyAsync(d_pi, pi.data(), total * 4, cudaMemcpyHostToDevice, c->stream));
            SPK_CUDA(c, cudaMemcpyAsync(d_pj, pj.data(), total * 4, cudaMemcpyHostToDevice, c->stream));
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_THR0], c->stream));
        } else {
            uint32_t half_bits = 1; while (half_bits < 32 && (1ull << (2 * half_bits)) < possible) ++half_bits;
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_THR0], c->stream));
            sample_pairs_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(d_pi, d_pj, tp, nit, rank, world, L, possible, half_bits, seed);
            c->tm.kernel_launches += 1;
        }
        // all pairs through the class-bucketed cooperative refine kernels (the order inside an iteration's segment is free: it is
        // sorted next); the thread-per-pair kernel only when the weights are not groupable
        r = mi_pairs_segmented(c, d_pi, d_pj, total, tp, nit, pc, d_mi, d_work);
        if (r == SPK_E_INVALID && c->g_state != 1) r = mi_pairs_exact(c, d_pi, d_pj, total, pc, d_mi);
        if (r) return r;
        SPK_CUDA(c, cudaMemcpyAsync(d_off, h_off.data(), (size_t)(nit + 1) * 4, cudaMemcpyHostToDevice, c->stream));
        segment_ids_kernel<<<(unsigned)((total + 255) / 256), 256, 0, c->stream>>>(d_seg, total, tp);
        SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(d_tmp, tmpb, (const double*)d_mi, d_byval, (const uint32_t*)d_seg, d_seg2, (int)total, 0, 64, c->stream));
        SPK_CUDA(c, cub::DeviceRadixSort::SortPairs(d_tmp, tmpb, (const uint32_t*)d_seg2, d_seg, (const double*)d_byval, d_sorted, (int)total, 0, (int)seg_bits, c->stream));
        // this rank's iterations sit at rank, rank + world, ...: write them to their global slots
        std::vector<double*> dummy;
        for (uint32_t k = 0; k < nit; ++k)
            pick_kth_kernel<<<1, 32, 0, c->stream>>>(d_sorted + (size_t)k * tp, tp, idx, 1, d_thr + my_its[k]);
        c->tm.kernel_launches += 9 + nit;
        SPK_CUDA(c, cudaEventRecord(c->ev[EV_THR1], c->stream));
        c->tm_pending |= TMP_THR;
    }
    r = comm_allreduce_sum_f64(c, d_thr, iterations); if (r) return r;
    std::vector<double> thresholds(iterations);
    SPK_CUDA(c, cudaMemcpyAsync(thresholds.data(), d_thr, (size_t)iterations * 8, cudaMemcpyDeviceToHost, c->stream));
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    resolve_timings(c);
    const size_t med = thresholds.size() / 2 - (thresholds.size() % 2 ? 0 : 1);       // SpydrPick.hpp:338-341
    std::nth_element(thresholds.begin(), thresholds.begin() + med, thresholds.end());
    *threshold = thresholds[med];
    return SPK_OK;
}

int spk_determine_mi_threshold(spk_ctx* c, uint64_t n_values, uint64_t threshold_pairs, uint32_t iterations,
                               double pc, uint64_t seed, double* threshold) {
    return spk_determine_mi_threshold_ex(c, n_values, threshold_pairs, iterations, pc, seed, SPK_SAMPLER_REFERENCE, threshold);
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

static int ensure_scan_buffers(Ctx* c, uint64_t want_edges, uint64_t want_cand, bool exhaustive) {
    const uint64_t possible = (uint64_t)c->L * (c->L - 1) / 2;
    want_edges = std::min<uint64_t>(want_edges, possible + 1);
    want_cand = std::min<uint64_t>(want_cand, possible + 1);
    want_cand += (uint64_t)c->sm_count * 16 * 2 * 128;        // int8 engine: two partly used emission chunks per epilogue warp
    if (exhaustive) want_cand = std::max<uint64_t>(std::min<uint64_t>(possible + 1, 1ull << 24), (uint64_t)c->L);
    int r;
    if (c->edges_cap < want_edges) {
        cudaStreamSynchronize(c->stream);
        if ((r = dev_alloc(c, c->d_edges, want_edges * sizeof(spk_edge), "edges"))) return r;
        if ((r = dev_alloc(c, c->d_wo, want_edges * sizeof(spk_edge), "edges_wo_gaps"))) return r;
        c->edges_cap = c->wo_cap = want_edges;
    }
    if (c->cand_cap < want_cand) {
        cudaStreamSynchronize(c->stream);
        if ((r = dev_alloc(c, c->d_cand, want_cand * sizeof(Candidate), "candidates"))) return r;
        if ((r = dev_alloc(c, c->d_refined, want_cand * sizeof(spk_edge), "refined"))) return r;
        if ((r = dev_alloc(c, c->d_refined_wo, want_cand * 8, "refined_wo"))) return r;
        c->cand_cap = c->refined_cap = want_cand;
    }
    if (!c->d_audit) {                                        // audit sample of the filter: 1M pairs (16 MB) is plenty
        const uint64_t cap = 1u << 20;
        if (cudaMalloc(&c->d_audit, cap * sizeof(Candidate)) == cudaSuccess) c->audit_cap = cap; else { cudaGetLastError(); c->d_audit = nullptr; c->audit_cap = 0; }
    }
    return SPK_OK;
}

// pinned result mirrors: D2H at full PCIe rate, no page-fault pass over a fresh std::vector
static int grow_pinned(Ctx* c, spk_edge*& buf, uint64_t& cap, uint64_t want) {
    if (want <= cap) return SPK_OK;
    if (buf) { cudaFreeHost(buf); buf = nullptr; cap = 0; }
    const uint64_t newcap = want + want / 4 + 1024;
    if (cudaHostAlloc((void**)&buf, newcap * sizeof(spk_edge), cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError(); return fail(c, SPK_E_NOMEM, "pinned host allocation failed (%llu edges)", (unsigned long long)newcap);
    }
    cap = newcap; return SPK_OK;
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
    // the library and a run's measured error (over the candidates AND over the audit sample of non-candidates) exceeds band/2, or
    // the audit finds a pair the filter would have lost, the scan is repeated with a wider band (never silently accepted).
    double band = p.band > 0.0 ? p.band : (use_i8 ? (i8_wr == 2 ? 4e-5 : 2e-4) : (T == 1 ? 1e-2 : 2e-4));
    const bool band_auto = !(p.band > 0.0);
    const uint64_t circle = p.circle_size ? p.circle_size : c->L;
    // n_parts == 0 on a context that belongs to a communicator: the collective scan (this rank's partition + the exchange)
    const bool collective = (p.n_parts == 0 && c->world > 1);
    if (collective) { p.part_index = (uint32_t)c->rank; p.n_parts = (uint32_t)c->world; }
    if (p.n_parts == 0) p.n_parts = 1;
    if (p.part_index >= p.n_parts) return fail(c, SPK_E_INVALID, "mi_scan: part_index >= n_parts");
    if (c->dl_pending) { SPK_CUDA(c, cudaEventSynchronize(c->ev_dl)); c->dl_pending = false; }
    const uint64_t possible = (uint64_t)c->L * (c->L - 1) / 2;
    uint64_t want_edges = p.max_edges ? p.max_edges : 2 * std::min<uint64_t>(10000000ull, 100ull * c->L) + 65536;
    if (p.n_parts > 1 && !p.max_edges) want_edges = want_edges / p.n_parts + want_edges / (4 * p.n_parts) + 65536;   // a partition holds ~1/G of the edges
    uint64_t want_cand = want_edges + want_edges / 2 + (uint64_t)c->L * 64 / p.n_parts + (1u << 20);
    int r = ensure_scan_buffers(c, want_edges, want_cand, p.engine == 1); if (r) return r;

    unsigned long long h_counters[16] = {0};
    uint64_t ne = 0, nw = 0;
    std::vector<double> xcnt;                                // collective: {edges, wo, worst error, lost} of every rank
    bool first = true;
    for (int attempt = 0;; ++attempt) {
        if (attempt > 8) return fail(c, SPK_E_OVERFLOW, "mi_scan: result buffers still too small after %d rescans", attempt);
        // counters [11], [12] (size / overflow of the audit sample) belong to the seed pass: a rescan that reuses the seeds keeps them
        if (first) SPK_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 16 * sizeof(unsigned long long), c->stream));
        else { SPK_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 11 * sizeof(unsigned long long), c->stream));
               SPK_CUDA(c, cudaMemsetAsync(c->d_counters + 13, 0, 3 * sizeof(unsigned long long), c->stream)); }
        fill_u64_kernel<<<(c->L_pad + 255) / 256, 256, 0, c->stream>>>((unsigned long long*)c->d_colmax, f64_ordered(-DBL_MAX), c->L_pad);
        c->tm.kernel_launches += 1;
        c->tm.gemm_ms = c->tm.refine_ms = c->tm.compact_ms = c->tm.sort_ms = c->tm.seed_ms = c->tm.audit_ms = c->tm.exchange_ms = 0.f; c->tm.gemm_tiles = 0;
        bool audited = false;
        if (p.engine == 1) {
            r = mi_scan_exact(c, p, circle); if (r) return r;
        } else {
            if (use_i8) {
                // the refine's plane build only needs the weights: queued first, it runs on the GPU while the host prepares the filter
                if (first) { r = mi_refine_prepare(c, p.pseudocount); if (r) return r; }
                r = mi_i8_filter(c, p, band, circle, i8_wr, !first); if (r) return r;
                if (c->d_audit && c->audit_cap) {
                    // the audit sample the seed pass left (this attempt's, or the one a rescan reuses) joins the candidate list as marked entries
                    r = mi_audit_append(c, c->d_audit, c->d_counters + 11, c->audit_cap); if (r) return r;
                    audited = true;
                }
            } else {
                r = prep_column_stats(c, p.pseudocount); if (r) return r;
                r = prep_operands(c, T); if (r) return r;
                r = mi_gemm_filter(c, p, T, band, circle); if (r) return r;
            }
            first = false;
            SPK_CUDA(c, cudaMemcpyAsync(h_counters, c->d_counters, sizeof h_counters, cudaMemcpyDeviceToHost, c->stream));
            SPK_CUDA(c, cudaStreamSynchronize(c->stream));
            if (h_counters[4]) {
                // candidate buffer too small: the reservation counter kept counting, so the needed size is known -- grow and rescan
                // (the reference stores any number of edges; a low user threshold must not abort the run)
                const uint64_t need = h_counters[0] + h_counters[0] / 8 + (1u << 16);
                if (c->verbose) fprintf(stderr, "spydrpick_b200: %llu candidates > capacity %llu: rescanning with larger buffers\n", h_counters[0], (unsigned long long)c->cand_cap);
                if (need > possible + (uint64_t)c->sm_count * 16 * 2 * 128 + (1u << 16) && c->cand_cap >= possible) return fail(c, SPK_E_OVERFLOW, "candidate buffer overflow");
                r = ensure_scan_buffers(c, std::max<uint64_t>(c->edges_cap, need * 2 / 3), need, false); if (r) return r;
                continue;
            }
            uint64_t ncand = h_counters[0];                  // int8 engine: reserved slots, a few of them invalid-marked holes; + the audit entries
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_REF0], c->stream));
            r = mi_refine_candidates(c, ncand, p.pseudocount, p.threshold, &ncand); if (r) return r;
            res->n_candidates = ncand;
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_REF1], c->stream));
            r = mi_compact(c, ncand, p, circle, audited ? (float)(p.threshold - band) : __builtin_nanf("")); if (r) return r;
            SPK_CUDA(c, cudaEventRecord(c->ev[EV_CMP1], c->stream));
            c->tm_pending |= TMP_REFINE | TMP_COMPACT;
        }
        ordered_to_double_kernel<<<(c->L_pad + 255) / 256, 256, 0, c->stream>>>((unsigned long long*)c->d_colmax, c->L_pad);
        c->tm.kernel_launches += 1;
        SPK_CUDA(c, cudaMemcpyAsync(h_counters, c->d_counters, sizeof h_counters, cudaMemcpyDeviceToHost, c->stream));
        SPK_CUDA(c, cudaStreamSynchronize(c->stream));
        resolve_timings(c);
        if (h_counters[4]) {
            const uint64_t need = std::max(h_counters[1], h_counters[2]) + h_counters[1] / 8 + (1u << 16);
            if (c->verbose) fprintf(stderr, "spydrpick_b200: %llu edges > capacity %llu: rescanning with larger buffers\n", h_counters[1], (unsigned long long)c->edges_cap);
            if (c->edges_cap > possible) return fail(c, SPK_E_OVERFLOW, "edge buffer overflow");
            r = ensure_scan_buffers(c, need, std::max<uint64_t>(c->cand_cap, need + need / 2), p.engine == 1); if (r) return r;
            continue;
        }
        ne = h_counters[1]; nw = h_counters[2];
        res->max_filter_error = h_counters[3] ? f64_from_ordered(h_counters[3]) : 0.0;
        if (audited) {                                       // written by the compaction from the marked entries
            res->audit_pairs = h_counters[10]; res->audit_violations = h_counters[9];
            res->audit_max_error = h_counters[8] ? f64_from_ordered(h_counters[8]) : 0.0;
            res->n_candidates -= std::min<uint64_t>(res->n_candidates, h_counters[10]);
        }
        double worst = std::max(res->max_filter_error, res->audit_max_error);
        bool lost = res->audit_violations != 0;
        if (collective) {
            // ONE 32-byte-per-rank all-gather carries the band decision (taken by all ranks together: a rescan with a wider band
            // repeats the collective seed pass) and the list lengths the exchange below needs
            if (!c->d_xcnt && cudaMalloc(&c->d_xcnt, 4 * 64 * sizeof(double)) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "exchange counters"); }
            xcnt.assign(4 * (size_t)c->world, 0.0);
            const double mine[4] = {(double)ne, (double)nw, worst, lost ? 1.0 : 0.0};       // counts < 2^53: exact in a double
            SPK_CUDA(c, cudaMemcpyAsync((double*)c->d_xcnt + 4 * c->rank, mine, 32, cudaMemcpyHostToDevice, c->stream));
            r = comm_allgather_inplace(c, c->d_xcnt, 32); if (r) return r;
            SPK_CUDA(c, cudaMemcpyAsync(xcnt.data(), c->d_xcnt, (size_t)c->world * 32, cudaMemcpyDeviceToHost, c->stream));
            SPK_CUDA(c, cudaStreamSynchronize(c->stream));
            for (int q = 0; q < c->world; ++q) { worst = std::max(worst, xcnt[4 * q + 2]); lost = lost || xcnt[4 * q + 3] != 0.0; }
        }
        if (p.engine != 1 && (worst > 0.5 * band || lost)) {
            if (band_auto && attempt < 6) {
                const double wider = std::max(4.0 * worst, 4.0 * band);
                if (c->verbose) fprintf(stderr, "spydrpick_b200: filter error %.3g (audit: %.3g over %llu pairs, %llu lost) vs band %.3g: rescanning with band %.3g\n",
                                        res->max_filter_error, res->audit_max_error, (unsigned long long)res->audit_pairs, (unsigned long long)res->audit_violations, band, wider);
                band = wider;
                first = true;                                   // a wider band needs a new audit sample (its window moves with the band)
                continue;
            }
            return fail(c, SPK_E_PRECISION, "filter error %.3g (audit %.3g, %llu pairs lost) exceeds half the safety band %.3g: raise band (or use engine 3)",
                        res->max_filter_error, res->audit_max_error, (unsigned long long)res->audit_violations, band);
        }
        break;
    }
    res->band = band;
    c->d_edges_sorted_n = 0; c->d_edges_view = c->d_edges;
    if (p.sort_desc) { r = sort_edges_device(c, c->d_edges, ne); if (r) return r; }
    spk_edge* d_list = c->d_edges; spk_edge* d_wolist = c->d_wo;
    uint64_t ne_all = ne, nw_all = nw;
    if (collective) {
        // ---- the one exchange step before ARACNE (MI_solver::join, include/mi.hpp:355-361): edge lists concatenated, colmax max-reduced
        const int G = c->world;
        SPK_CUDA(c, cudaEventRecord(c->ev[EV_XCH0], c->stream));
        std::vector<unsigned long long> cnt(2 * G);
        for (int q = 0; q < G; ++q) { cnt[2 * q] = (unsigned long long)xcnt[4 * q]; cnt[2 * q + 1] = (unsigned long long)xcnt[4 * q + 1]; }
        r = comm_allreduce_max_f64(c, c->d_colmax, c->L); if (r) return r;
        std::vector<size_t> eb(G), eo(G), wb(G), wo(G);
        std::vector<uint64_t> runs(G);
        ne_all = 0; nw_all = 0;
        for (int q = 0; q < G; ++q) { eo[q] = ne_all * sizeof(spk_edge); eb[q] = cnt[2 * q] * sizeof(spk_edge); runs[q] = cnt[2 * q]; ne_all += cnt[2 * q];
                                      wo[q] = nw_all * sizeof(spk_edge); wb[q] = cnt[2 * q + 1] * sizeof(spk_edge); nw_all += cnt[2 * q + 1]; }
        if (ne_all > 0x7fffffffull) return fail(c, SPK_E_INVALID, "mi_scan: more than 2^31 edges in total");
        if (c->merge_cap < ne_all) {
            free_dev(c->d_merge); c->merge_cap = 0;
            const uint64_t cap = ne_all + ne_all / 8 + 4096;
            if (cudaMalloc(&c->d_merge, cap * sizeof(spk_edge)) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "merged edge list (%llu edges)", (unsigned long long)cap); }
            c->merge_cap = cap;
        }
        r = comm_allgatherv(c, c->d_edges, c->d_merge, eb.data(), eo.data()); if (r) return r;
        if (c->rank == 0 && c->wo_all_cap < nw_all) {
            free_dev(c->d_wo_all); c->wo_all_cap = 0;
            const uint64_t cap = nw_all + nw_all / 8 + 4096;
            if (cudaMalloc(&c->d_wo_all, cap * sizeof(spk_edge)) != cudaSuccess) { cudaGetLastError(); return fail(c, SPK_E_NOMEM, "gathered MI-without-gaps list"); }
            c->wo_all_cap = cap;
        }
        r = comm_gatherv_root(c, c->d_wo, c->d_wo_all, wb.data(), wo.data()); if (r) return r;
        // every rank sorted its own edges: the gathered list is G sorted runs -> G-1 merges give the list a full sort gives
        if (p.sort_desc) { r = merge_sorted_runs_device(c, c->d_merge, runs.data(), (uint32_t)G); if (r) return r; }
        SPK_CUDA(c, cudaEventRecord(c->ev[EV_XCH1], c->stream));
        c->tm_pending |= TMP_XCH;
        d_list = c->d_merge; d_wolist = (c->rank == 0) ? c->d_wo_all : nullptr;
        if (c->rank != 0) nw_all = 0;
        c->d_edges_view = c->d_merge;
    }
    if (p.sort_desc) c->d_edges_sorted_n = ne_all;
    // colmax is final: read back BEFORE the bulk downloads are queued (one D2H copy engine, in issue order)
    c->h_colmax.resize(c->L);
    SPK_CUDA(c, cudaMemcpyAsync(c->h_colmax.data(), c->d_colmax, (size_t)c->L * 8, cudaMemcpyDeviceToHost, c->stream));
    res->n_edges = ne_all; res->n_wo_gaps = nw_all;
    res->d_edges = d_list; res->d_edges_wo_gaps = d_wolist; res->d_colmax = c->d_colmax;
    c->host_edges_n = 0;
    const bool want_host = p.keep_on_device != 1 && (!collective || c->rank == 0);
    if (want_host) {
        if ((r = grow_pinned(c, c->hp_edges, c->hp_edges_cap, ne_all))) return r;
        if ((r = grow_pinned(c, c->hp_wo, c->hp_wo_cap, nw_all))) return r;
        // keep_on_device == 2: the copies run on a side stream and overlap whatever the caller launches next (ARACNE on the
        // device copy); the host pointers are returned now, their contents are valid after spk_wait_download().
        cudaStream_t cs = c->stream;
        if (p.keep_on_device == 2) {
            SPK_CUDA(c, cudaEventRecord(c->ev_ready, c->stream));
            SPK_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev_ready, 0));
            cs = c->copy_stream;
        }
        if (nw_all) SPK_CUDA(c, cudaMemcpyAsync(c->hp_wo, d_wolist, nw_all * sizeof(spk_edge), cudaMemcpyDeviceToHost, cs));
        if (ne_all) SPK_CUDA(c, cudaMemcpyAsync(c->hp_edges, d_list, ne_all * sizeof(spk_edge), cudaMemcpyDeviceToHost, cs));
        if (p.keep_on_device == 2) { SPK_CUDA(c, cudaEventRecord(c->ev_dl, cs)); c->dl_pending = true; }
        if (p.sort_desc) c->host_edges_n = ne_all;
    }
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    resolve_timings(c);
    res->colmax = c->h_colmax.data();
    if (want_host) { res->edges = c->hp_edges; res->edges_wo_gaps = c->hp_wo; }
    return SPK_OK;
}

int spk_sort_edges_device(spk_ctx* c, void* d_edges, uint64_t n) {
    if (!c || (!d_edges && n)) return SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = sort_edges_device(c, (spk_edge*)d_edges, n); if (r) return r;
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    resolve_timings(c);
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
    if (!r) resolve_timings(c);
    cudaFree(d);
    return r;
}

int spk_merge_sorted_runs_device(spk_ctx* c, void* d_edges, const uint64_t* run_counts, uint32_t n_runs) {
    if (!c || !run_counts) return SPK_E_INVALID;
    uint64_t n = 0;
    for (uint32_t r = 0; r < n_runs; ++r) n += run_counts[r];
    if (n && !d_edges) return SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r = merge_sorted_runs_device(c, (spk_edge*)d_edges, run_counts, n_runs); if (r) return r;
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    return SPK_OK;
}

// ARACNE on a device list. Collective (context in a communicator, n_parts == 0): every rank evaluates its range of the edge
// list, the flag shards are exchanged in place (grouped broadcast of each rank's range), every rank ends with all flags.
// collective ARACNE: every rank evaluates the edges owned by ITS nodes (u = rank mod world; an edge is owned by its larger-degree
// endpoint) into zero-initialised flags, and one all-reduce(max) of E bytes gives every rank the whole vector
static int aracne_collective(spk_ctx* c, const spk_edge* d_edges, uint64_t n, double eps, uint8_t* d_flags) {
    SPK_CUDA(c, cudaMemsetAsync(d_flags, 0, n, c->stream));
    int r = aracne_device(c, d_edges, n, eps, (uint32_t)c->rank, (uint32_t)c->world, d_flags, true); if (r) return r;
    return comm_allreduce_max_u8(c, d_flags, n);
}

int spk_aracne_device(spk_ctx* c, const void* d_edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, void* d_flags) {
    if (!c || (n && (!d_edges || !d_flags))) return SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    int r;
    if (nparts == 0 && c->world > 1) {
        r = aracne_collective(c, (const spk_edge*)d_edges, n, eps, (uint8_t*)d_flags); if (r) return r;
    } else {
        if (nparts == 0) nparts = 1;
        if (part >= nparts) return fail(c, SPK_E_INVALID, "aracne: part_index >= n_parts");
        r = aracne_device(c, (const spk_edge*)d_edges, n, eps, part, nparts, (uint8_t*)d_flags); if (r) return r;
    }
    SPK_CUDA(c, cudaStreamSynchronize(c->stream));
    resolve_timings(c);
    return SPK_OK;
}

int spk_aracne(spk_ctx* c, const spk_edge* edges, uint64_t n, double eps, uint32_t part, uint32_t nparts, uint8_t* flags) {
    if (!c) return SPK_E_INVALID;
    const bool collective = (nparts == 0 && c->world > 1);
    // collective: every rank calls with the same n (the scan's n_edges); ranks other than 0 hold no host copy and may pass
    // edges = flags = NULL -- the flags are delivered to rank 0's host buffer
    if (!collective && n && (!edges || !flags)) return SPK_E_INVALID;
    if (n == 0) return SPK_OK;
    if (nparts == 0) nparts = 1;
    if (!collective && part >= nparts) return fail(c, SPK_E_INVALID, "aracne: part_index >= n_parts");
    SPK_CUDA(c, cudaSetDevice(c->device));
    const bool own_list = c->d_edges_view && c->d_edges_sorted_n == n && (collective ? true : (edges == c->hp_edges && n == c->host_edges_n));
    if (collective && !own_list) return fail(c, SPK_E_INVALID, "aracne: the collective form runs on the edge list of the last collective scan");
    if (own_list) {
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
        int r0 = (nparts <= 1 || collective) ? cuda_check(c, cudaMemsetAsync(f0, 0, n, c->stream), "clear flags")
                                             : cuda_check(c, cudaMemcpyAsync(f0, flags, n, cudaMemcpyHostToDevice, c->stream), "H2D flags");
        if (!r0) r0 = collective ? aracne_collective(c, c->d_edges_view, n, eps, f0) : aracne_device(c, c->d_edges_view, n, eps, part, nparts, f0);
        if (!r0 && flags) r0 = cuda_check(c, cudaMemcpyAsync(flags, f0, n, cudaMemcpyDeviceToHost, c->stream), "D2H flags");
        if (!r0) r0 = cuda_check(c, cudaStreamSynchronize(c->stream), "sync");
        if (!r0) resolve_timings(c);
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
    if (!r) resolve_timings(c);
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
    if (c->tm_pending) { cudaSetDevice(c->device); cudaStreamSynchronize(c->stream); resolve_timings(c); }
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

int spk_debug_i8_peak(spk_ctx* c, uint32_t kblocks, uint32_t reps, double* tops, float* ms) {
    if (!c) return SPK_E_INVALID;
    SPK_CUDA(c, cudaSetDevice(c->device));
    return i8_tensor_peak(c, kblocks, reps, ms, tops);
}

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
