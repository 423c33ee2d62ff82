// spk_prep.cu -- packing kernels: column-major bytes, bit planes, presence masks,
// fp64 column statistics and the bf16 one-hot GEMM operands.
// All are HBM-streaming passes over the alignment (n*L bytes); none is on the O(L^2) path.
#include <cstring>
#include "spk_internal.h"

namespace spk {

// ---- [n][L] -> [L_pad][n_pad] transpose (32x32 byte tiles through shared memory) ---------
__global__ void transpose_colmajor_kernel(const uint8_t* __restrict__ states, uint8_t* __restrict__ colmajor,
                                          uint32_t n, uint32_t L, uint32_t n_pad) {
    __shared__ uint8_t tile[32][33];
    const uint32_t c0 = blockIdx.x * 32, s0 = blockIdx.y * 32;
    for (uint32_t r = threadIdx.y; r < 32; r += blockDim.y) {
        const uint32_t s = s0 + r, c = c0 + threadIdx.x;
        tile[r][threadIdx.x] = (s < n && c < L) ? states[(size_t)s * L + c] : (uint8_t)GAP;
    }
    __syncthreads();
    for (uint32_t r = threadIdx.y; r < 32; r += blockDim.y) {
        const uint32_t c = c0 + r, s = s0 + threadIdx.x;
        if (c < L && s < n) colmajor[(size_t)c * n_pad + s] = tile[threadIdx.x][r];
    }
}

// ---- bit planes for K1: planes[p][s][w] bit b = bit p of state(s, 32w+b); pad bits = 0 ----
__global__ void pack_planes_kernel(const uint8_t* __restrict__ states, uint32_t* __restrict__ planes,
                                   uint32_t n, uint32_t L, uint32_t W32) {
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

// ---- presence masks: one warp per column (Alignment_impl_block_compressed_storage.hpp:1089-1128)
__global__ void masks_kernel(const uint8_t* __restrict__ colmajor, uint8_t* __restrict__ masks,
                             uint32_t n, uint32_t L, uint32_t n_pad) {
    const uint32_t c = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (c >= L) return;
    const uint8_t* col = colmajor + (size_t)c * n_pad;
    uint32_t m = 0;
    for (uint32_t s = threadIdx.x % 32; s < n; s += 32) m |= 1u << col[s];
    for (int o = 16; o; o >>= 1) m |= __shfl_xor_sync(0xffffffffu, m, o);
    if ((threadIdx.x % 32) == 0) masks[c] = (uint8_t)(m & 31u);
}

// ---- fp64 weighted marginals + epilogue statistics: one warp per column --------------------
__global__ void column_stats_kernel(const uint8_t* __restrict__ colmajor, const double* __restrict__ w,
                                    const uint8_t* __restrict__ masks, const uint64_t* __restrict__ pos,
                                    double* __restrict__ marg, ColStat* __restrict__ stat,
                                    uint32_t n, uint32_t L, uint32_t L_pad, uint32_t n_pad, double w_total, double pc) {
    const uint32_t c = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    const uint32_t lane = threadIdx.x % 32;
    if (c >= L_pad) return;
    double m[NSTATES] = {0, 0, 0, 0, 0};
    if (c < L) {
        const uint8_t* col = colmajor + (size_t)c * n_pad;
        for (uint32_t s = lane; s < n; s += 32) {
            const uint32_t x = col[s]; const double ws = w[s];
#pragma unroll
            for (int a = 0; a < NSTATES; ++a) m[a] += (x == (uint32_t)a) ? ws : 0.0;
        }
#pragma unroll
        for (int a = 0; a < NSTATES; ++a)
            for (int o = 16; o; o >>= 1) m[a] += __shfl_xor_sync(0xffffffffu, m[a], o);
    }
    if (lane == 0) {
        ColStat st; memset(&st, 0, sizeof st);
        const uint32_t mask = (c < L) ? masks[c] : 0u;
        const uint32_t np = __popc(mask);
        for (int a = 0; a < NSTATES; ++a) { marg[(size_t)c * NSTATES + a] = m[a]; st.m[a] = (float)m[a]; }
        for (int k = 1; k <= NSTATES; ++k) {
            const double Z = w_total + pc * (double)np * (double)k;
            double h = 0.0;
            for (int a = 0; a < NSTATES; ++a) if (mask & (1u << a)) { const double p = (m[a] + pc * (double)k) / Z; h += p * log(p); }
            st.h[k - 1] = (float)h;
        }
        st.mask = mask; st.npres = np;
        const uint64_t p = (c < L) ? pos[c] : 0ull;
        st.pos_lo = (uint32_t)p; st.pos_hi = (uint32_t)(p >> 32);
        stat[c] = st;
    }
}

// ---- weight split into bf16 terms: w = t0 + t1 + t2 (+ O(2^-24 w)) -------------------------
__global__ void split_weights_kernel(const double* __restrict__ w, __nv_bfloat16* __restrict__ wsplit,
                                     uint32_t n, uint32_t n_pad, int n_terms) {
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_pad) return;
    double r = (s < n) ? w[s] : 0.0;
    for (int t = 0; t < n_terms; ++t) {
        const __nv_bfloat16 h = __double2bfloat16(r);
        wsplit[(size_t)t * n_pad + s] = h;
        r -= (double)__bfloat162float(h);
    }
}

// ---- one-hot expansion: U[(4c+a)][s] = [x==a], V_t[(4c+a)][s] = [x==a] * w_t[s] -------------
// one warp per column; each lane converts 8 consecutive samples per iteration (16-byte stores).
__global__ void expand_operands_kernel(const uint8_t* __restrict__ colmajor, const __nv_bfloat16* __restrict__ wsplit,
                                       __nv_bfloat16* __restrict__ U, __nv_bfloat16* __restrict__ V,
                                       uint32_t L_pad, uint32_t n_pad, int n_terms) {
    const uint32_t c = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (c >= L_pad) return;
    const uint32_t lane = threadIdx.x % 32;
    const uint8_t* col = colmajor + (size_t)c * n_pad;
    const size_t op_elems = (size_t)4 * L_pad * n_pad;
    const __nv_bfloat16 one = __float2bfloat16(1.0f), zero = __float2bfloat16(0.0f);
    for (uint32_t s0 = lane * 8; s0 < n_pad; s0 += 256) {
        const uint2 raw = *reinterpret_cast<const uint2*>(col + s0);
        uint8_t x[8];
        memcpy(x, &raw, 8);
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            __align__(16) __nv_bfloat16 u[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) u[k] = (x[k] == a) ? one : zero;
            const size_t off = ((size_t)(4 * c + a)) * n_pad + s0;
            *reinterpret_cast<uint4*>(U + off) = *reinterpret_cast<const uint4*>(u);
            for (int t = 0; t < n_terms; ++t) {
                __align__(16) __nv_bfloat16 wv[8];
                *reinterpret_cast<uint4*>(wv) = *reinterpret_cast<const uint4*>(wsplit + (size_t)t * n_pad + s0);
                __align__(16) __nv_bfloat16 v[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] = (x[k] == a) ? wv[k] : zero;
                *reinterpret_cast<uint4*>(V + (size_t)t * op_elems + off) = *reinterpret_cast<const uint4*>(v);
            }
        }
    }
}

// =============================================================================================
int prep_alignment(Ctx* c) {
    const uint32_t n = c->n, L = c->L;
    SPK_CUDA(c, cudaMemsetAsync(c->d_colmajor, GAP, (size_t)c->L_pad * c->n_pad, c->stream));
    dim3 tb(32, 8), tg((L + 31) / 32, (n + 31) / 32);
    transpose_colmajor_kernel<<<tg, tb, 0, c->stream>>>(c->d_states, c->d_colmajor, n, L, c->n_pad);
    const size_t nw = (size_t)n * c->W32;
    pack_planes_kernel<<<(unsigned)((nw + 255) / 256), 256, 0, c->stream>>>(c->d_states, c->d_planes, n, L, c->W32);
    SPK_CUDA(c, cudaMemsetAsync(c->d_masks, 0, c->L_pad, c->stream));
    masks_kernel<<<(L + 7) / 8, 256, 0, c->stream>>>(c->d_colmajor, c->d_masks, n, L, c->n_pad);
    c->tm.kernel_launches += 3;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

int prep_column_stats(Ctx* c, double pc) {
    if (c->stats_valid && c->stats_pc == pc) return SPK_OK;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_TMP0], c->stream));
    column_stats_kernel<<<(c->L_pad + 7) / 8, 256, 0, c->stream>>>(c->d_colmajor, c->d_w, c->d_masks, c->d_pos,
                                                                   c->d_marg, c->d_colstat, c->n, c->L, c->L_pad,
                                                                   c->n_pad, c->w_total, pc);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_TMP1], c->stream));
    SPK_CUDA(c, cudaEventSynchronize(c->ev[EV_TMP1]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.stats_ms, c->ev[EV_TMP0], c->ev[EV_TMP1]));
    SPK_CUDA(c, cudaGetLastError());
    c->stats_valid = true; c->stats_pc = pc;
    return SPK_OK;
}

int prep_operands(Ctx* c, int n_terms) {
    if (c->operands_terms == n_terms) return SPK_OK;
    const size_t op_elems = (size_t)4 * c->L_pad * c->n_pad;
    if (!c->d_U) {
        if (cudaMalloc(&c->d_U, op_elems * sizeof(__nv_bfloat16)) != cudaSuccess)
            return fail(c, SPK_E_NOMEM, "cannot allocate one-hot operand U (%zu bytes)", op_elems * 2);
    }
    if (c->operands_cap_terms < (size_t)n_terms) {
        if (c->d_V) { cudaFree(c->d_V); c->d_V = nullptr; }
        if (cudaMalloc(&c->d_V, op_elems * sizeof(__nv_bfloat16) * n_terms) != cudaSuccess)
            return fail(c, SPK_E_NOMEM, "cannot allocate weighted operand V (%zu bytes)", op_elems * 2 * n_terms);
        c->operands_cap_terms = n_terms;
    }
    __nv_bfloat16* wsplit = (__nv_bfloat16*)scratch(c, (size_t)3 * c->n_pad * sizeof(__nv_bfloat16));
    if (!wsplit) return fail(c, SPK_E_NOMEM, "scratch allocation failed");
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_TMP0], c->stream));
    split_weights_kernel<<<(c->n_pad + 255) / 256, 256, 0, c->stream>>>(c->d_w, wsplit, c->n, c->n_pad, n_terms);
    expand_operands_kernel<<<(c->L_pad + 7) / 8, 256, 0, c->stream>>>(c->d_colmajor, wsplit, c->d_U, c->d_V,
                                                                      c->L_pad, c->n_pad, n_terms);
    c->tm.kernel_launches += 2;
    SPK_CUDA(c, cudaEventRecord(c->ev[EV_TMP1], c->stream));
    SPK_CUDA(c, cudaEventSynchronize(c->ev[EV_TMP1]));
    SPK_CUDA(c, cudaEventElapsedTime(&c->tm.expand_ms, c->ev[EV_TMP0], c->ev[EV_TMP1]));
    SPK_CUDA(c, cudaGetLastError());
    c->operands_terms = n_terms;
    return SPK_OK;
}

// ---- column filter pre-pass -------------------------------------------------------------------
// Replaces the frequency scan of apegrunt::Alignment_filter (apegrunt/Alignment_filter.hpp:250-339): unweighted counts of
// the five states per column over the RAW alignment, O(n L) bytes read once (HBM-bound: n*L bytes in, 20 L bytes out).
// One thread per column (a warp reads 32 consecutive bytes of a row: whole sectors), rows dealt to blockIdx.y in chunks;
// every thread adds its five partial counts to the global table.
__global__ void __launch_bounds__(256)
column_counts_kernel(const uint8_t* __restrict__ states, uint32_t n, unsigned long long L, uint32_t rows_per_chunk,
                     uint32_t* __restrict__ counts) {
    const unsigned long long c = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= L) return;
    const uint32_t r0 = blockIdx.y * rows_per_chunk, r1 = min(n, r0 + rows_per_chunk);
    uint32_t k0 = 0, k1 = 0, k2 = 0, k3 = 0, k4 = 0;
    const uint8_t* ptr = states + (size_t)r0 * L + c;
#pragma unroll 4
    for (uint32_t s = r0; s < r1; ++s, ptr += L) {
        const uint32_t x = *ptr;
        k0 += (x == 0u); k1 += (x == 1u); k2 += (x == 2u); k3 += (x == 3u); k4 += (x >= 4u);
    }
    uint32_t* out = counts + c * NSTATES;
    if (k0) atomicAdd(out + 0, k0);
    if (k1) atomicAdd(out + 1, k1);
    if (k2) atomicAdd(out + 2, k2);
    if (k3) atomicAdd(out + 3, k3);
    if (k4) atomicAdd(out + 4, k4);
}

// filtered[s][k] = raw[s][idx[k]] (idx ascending: reads nearly coalesced, writes coalesced)
__global__ void __launch_bounds__(256)
gather_columns_kernel(const uint8_t* __restrict__ raw, unsigned long long L, const uint32_t* __restrict__ idx, uint32_t Lk,
                      uint8_t* __restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= Lk) return;
    const size_t s = blockIdx.y;
    out[s * Lk + k] = raw[s * L + idx[k]];
}

int column_state_counts_device(Ctx* c, const uint8_t* d_raw, uint32_t n, uint64_t L, uint32_t* d_counts) {
    SPK_CUDA(c, cudaMemsetAsync(d_counts, 0, (size_t)L * NSTATES * 4, c->stream));
    const uint32_t rows_per_chunk = 128;
    const unsigned gx = (unsigned)((L + 255) / 256), gy = (n + rows_per_chunk - 1) / rows_per_chunk;
    column_counts_kernel<<<dim3(gx, gy), 256, 0, c->stream>>>(d_raw, n, (unsigned long long)L, rows_per_chunk, d_counts);
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

int gather_columns_device(Ctx* c, const uint8_t* d_raw, uint32_t n, uint64_t L, const uint32_t* d_idx, uint32_t Lk, uint8_t* d_out) {
    if (n > 65535u) {                                    // gridDim.y limit: rows in slabs
        for (uint32_t r0 = 0; r0 < n; r0 += 65535u) {
            const uint32_t rows = min(65535u, n - r0);
            gather_columns_kernel<<<dim3((Lk + 255) / 256, rows), 256, 0, c->stream>>>(d_raw + (size_t)r0 * L, (unsigned long long)L, d_idx, Lk, d_out + (size_t)r0 * Lk);
        }
    } else {
        gather_columns_kernel<<<dim3((Lk + 255) / 256, n), 256, 0, c->stream>>>(d_raw, (unsigned long long)L, d_idx, Lk, d_out);
    }
    c->tm.kernel_launches += 1;
    SPK_CUDA(c, cudaGetLastError());
    return SPK_OK;
}

} // namespace spk
