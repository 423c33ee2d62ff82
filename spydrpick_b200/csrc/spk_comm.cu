// spk_comm.cu -- multi-GPU plumbing behind the C-ABI: one context per device, SPMD (every rank's context makes the same calls).
//
// What is exchanged (the reference's counterpart is MI_solver::join, include/mi.hpp:355-361: edge lists are
// concatenated and the per-column maxima max-reduced when TBB joins the bodies of a split range):
//   K1      partial neighbour counts          ncclAllReduce(sum, u64[n])
//   scan    colmax seed pass (fp32, sharded)  ncclAllReduce(max, u32[Lperm])  (ordered-uint encoding)
//           refine bit planes (sharded build) ncclAllGather
//           per-rank edge lists               counts all-gathered, then ONE grouped ncclBroadcast per rank straight into the
//                                             merge buffer at its prefix offset (no padding, no concatenation pass)
//           MI-without-gaps lists             grouped ncclSend / ncclRecv to rank 0 (only the writer needs them)
//           per-column maxima                 ncclAllReduce(max, f64[L])
//   ARACNE  direct-edge flags (node shards)   ncclAllReduce(max, u8[E]) over zero-initialised flags
//   upload  alignment rows                    each rank uploads n/G rows over its own PCIe link, ncclAllGather over NVLink
//
// NCCL is loaded with dlopen (libnccl.so.2): inside a Python process that is torch's bundled copy (already mapped), for the
// C++ host the system library. Only the handful of entry points below are used; their signatures are stable across NCCL 2.x.
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>
#include "spk_internal.h"

namespace spk {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { NCCL_CHAR = 0, NCCL_UINT8 = 1, NCCL_UINT32 = 3, NCCL_UINT64 = 5, NCCL_FLOAT64 = 8 };
enum { NCCL_SUM = 0, NCCL_MAX = 2 };

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char* (*GetErrorString)(ncclResult_t);
    ncclResult_t (*GetVersion)(int*);
    void* handle = nullptr;
    std::string error;
};

static NcclApi* nccl_api(std::string* err) {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* env = getenv("SPYDRPICK_B200_NCCL");
        const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) {
            if (!nm || !*nm) continue;
            api.handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
            if (api.handle) break;
        }
        if (!api.handle) { api.error = std::string("cannot load libnccl.so.2 (set SPYDRPICK_B200_NCCL): ") + (dlerror() ? dlerror() : ""); return; }
        auto sym = [&](const char* s) { void* p = dlsym(api.handle, s); if (!p && api.error.empty()) api.error = std::string("NCCL symbol missing: ") + s; return p; };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
        api.Broadcast = (decltype(api.Broadcast))sym("ncclBroadcast");
        api.Send = (decltype(api.Send))sym("ncclSend");
        api.Recv = (decltype(api.Recv))sym("ncclRecv");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
    });
    if (!api.error.empty()) { if (err) *err = api.error; return nullptr; }
    return &api;
}

static int nccl_check(Ctx* c, ncclResult_t r, const char* what) {
    if (r == 0) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    return fail(c, SPK_E_CUDA, "NCCL %s failed: %s", what, a && a->GetErrorString ? a->GetErrorString(r) : "?");
}
#define SPK_NCCL(c, call, what) do { int _r = nccl_check((c), (call), (what)); if (_r) return _r; } while (0)

int comm_allreduce_sum_u64(Ctx* c, unsigned long long* d, size_t n) {
    if (c->world <= 1) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->AllReduce(d, d, n, NCCL_UINT64, NCCL_SUM, (ncclComm_t)c->comm, c->stream), "allreduce(sum,u64)");
    return SPK_OK;
}
int comm_allreduce_max_u32(Ctx* c, uint32_t* d, size_t n) {
    if (c->world <= 1) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->AllReduce(d, d, n, NCCL_UINT32, NCCL_MAX, (ncclComm_t)c->comm, c->stream), "allreduce(max,u32)");
    return SPK_OK;
}
int comm_allreduce_max_u8(Ctx* c, uint8_t* d, size_t n) {
    if (c->world <= 1 || n == 0) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->AllReduce(d, d, n, NCCL_UINT8, NCCL_MAX, (ncclComm_t)c->comm, c->stream), "allreduce(max,u8)");
    return SPK_OK;
}
int comm_allreduce_max_f64(Ctx* c, double* d, size_t n) {
    if (c->world <= 1) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->AllReduce(d, d, n, NCCL_FLOAT64, NCCL_MAX, (ncclComm_t)c->comm, c->stream), "allreduce(max,f64)");
    return SPK_OK;
}
int comm_allreduce_sum_f64(Ctx* c, double* d, size_t n) {
    if (c->world <= 1) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->AllReduce(d, d, n, NCCL_FLOAT64, NCCL_SUM, (ncclComm_t)c->comm, c->stream), "allreduce(sum,f64)");
    return SPK_OK;
}
// every rank contributes bytes_per_rank bytes at d_buf + rank * bytes_per_rank (in place)
int comm_allgather_inplace(Ctx* c, void* d_buf, size_t bytes_per_rank) {
    if (c->world <= 1 || bytes_per_rank == 0) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->AllGather((const char*)d_buf + (size_t)c->rank * bytes_per_rank, d_buf, bytes_per_rank, NCCL_CHAR, (ncclComm_t)c->comm, c->stream), "allgather");
    return SPK_OK;
}
// variable-length all-gather: rank r's `bytes[r]` bytes (on rank r at d_mine) land at d_out + off[r] on every rank
int comm_allgatherv(Ctx* c, const void* d_mine, void* d_out, const size_t* bytes, const size_t* off) {
    if (c->world <= 1) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->GroupStart(), "group start");
    for (int r = 0; r < c->world; ++r) {
        if (bytes[r] == 0) continue;
        const void* src = (r == c->rank) ? d_mine : (const void*)((const char*)d_out + off[r]);
        SPK_NCCL(c, a->Broadcast(src, (char*)d_out + off[r], bytes[r], NCCL_CHAR, r, (ncclComm_t)c->comm, c->stream), "broadcast");
    }
    SPK_NCCL(c, a->GroupEnd(), "group end");
    return SPK_OK;
}
// variable-length gather to rank 0
int comm_gatherv_root(Ctx* c, const void* d_mine, void* d_out, const size_t* bytes, const size_t* off) {
    if (c->world <= 1) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->GroupStart(), "group start");
    if (c->rank == 0) {
        for (int r = 1; r < c->world; ++r)
            if (bytes[r]) SPK_NCCL(c, a->Recv((char*)d_out + off[r], bytes[r], NCCL_CHAR, r, (ncclComm_t)c->comm, c->stream), "recv");
    } else if (bytes[c->rank]) {
        SPK_NCCL(c, a->Send(d_mine, bytes[c->rank], NCCL_CHAR, 0, (ncclComm_t)c->comm, c->stream), "send");
    }
    SPK_NCCL(c, a->GroupEnd(), "group end");
    if (c->rank == 0 && bytes[0] && d_mine != (const char*)d_out + off[0])
        SPK_CUDA(c, cudaMemcpyAsync((char*)d_out + off[0], d_mine, bytes[0], cudaMemcpyDeviceToDevice, c->stream));
    return SPK_OK;
}
int comm_broadcast(Ctx* c, void* d_buf, size_t bytes, int root) {
    if (c->world <= 1 || bytes == 0) return SPK_OK;
    NcclApi* a = nccl_api(nullptr);
    SPK_NCCL(c, a->Broadcast(d_buf, d_buf, bytes, NCCL_CHAR, root, (ncclComm_t)c->comm, c->stream), "broadcast");
    return SPK_OK;
}

void comm_free(Ctx* c) {
    if (c->comm) { NcclApi* a = nccl_api(nullptr); if (a) a->CommDestroy((ncclComm_t)c->comm); c->comm = nullptr; }
    c->rank = 0; c->world = 1;
}

} // namespace spk

using namespace spk;

extern "C" {

int spk_comm_unique_id(void* id128) {
    if (!id128) return SPK_E_INVALID;
    std::string err;
    NcclApi* a = nccl_api(&err);
    if (!a) return fail(nullptr, SPK_E_CUDA, "%s", err.c_str());
    ncclUniqueId id;
    if (a->GetUniqueId(&id) != 0) return fail(nullptr, SPK_E_CUDA, "ncclGetUniqueId failed");
    memcpy(id128, &id, sizeof id);
    return SPK_OK;
}

int spk_comm_init_rank(spk_ctx* c, const void* id128, int rank, int world) {
    if (!c || !id128 || world < 1 || rank < 0 || rank >= world) return c ? fail(c, SPK_E_INVALID, "comm_init_rank: bad arguments") : SPK_E_INVALID;
    comm_free(c);
    if (world == 1) return SPK_OK;
    std::string err;
    NcclApi* a = nccl_api(&err);
    if (!a) return fail(c, SPK_E_CUDA, "%s", err.c_str());
    SPK_CUDA(c, cudaSetDevice(c->device));
    ncclUniqueId id; memcpy(&id, id128, sizeof id);
    ncclComm_t comm = nullptr;
    SPK_NCCL(c, a->CommInitRank(&comm, world, id, rank), "ncclCommInitRank");
    c->comm = comm; c->rank = rank; c->world = world;
    return SPK_OK;
}

int spk_comm_init_all(spk_ctx** ctxs, int n) {
    if (!ctxs || n < 1) return SPK_E_INVALID;
    for (int r = 0; r < n; ++r) if (!ctxs[r]) return SPK_E_INVALID;
    for (int r = 0; r < n; ++r) comm_free(ctxs[r]);
    if (n == 1) return SPK_OK;
    std::string err;
    NcclApi* a = nccl_api(&err);
    if (!a) return fail(ctxs[0], SPK_E_CUDA, "%s", err.c_str());
    std::vector<int> devs(n); std::vector<ncclComm_t> comms(n, nullptr);
    for (int r = 0; r < n; ++r) {
        devs[r] = ctxs[r]->device;
        for (int q = 0; q < r; ++q) if (devs[q] == devs[r]) return fail(ctxs[0], SPK_E_INVALID, "comm_init_all: contexts %d and %d share device %d", q, r, devs[r]);
    }
    ncclResult_t rc = a->CommInitAll(comms.data(), n, devs.data());
    if (rc != 0) return fail(ctxs[0], SPK_E_CUDA, "ncclCommInitAll failed: %s", a->GetErrorString(rc));
    for (int r = 0; r < n; ++r) { ctxs[r]->comm = comms[r]; ctxs[r]->rank = r; ctxs[r]->world = n; }
    return SPK_OK;
}

int spk_comm_rank(const spk_ctx* c, int* rank, int* world) {
    if (!c) return SPK_E_INVALID;
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    return SPK_OK;
}

} // extern "C"
