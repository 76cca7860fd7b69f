// =====================================================================================
// cvx_comm.cu -- the multi-GPU exchange of the edge-sketch path behind the C-ABI (include/curvix_b200.h, "multi-GPU").
//
// The path shards with no data-path collective (every view-1 edge / hypothesis view pair is independent,
// EdgeSketch_Core.cpp:209, :515-523); the only exchange is the gather of accepted rows + 3-D edges + tangents to the rank that
// runs the EdgeMapping insertions (EdgeSketch_Core.cpp:682-726).  Rounds: every rank stages the results of one unit on its
// context stream, a 64-byte record per rank is all-gathered (counts + what the unit was), then the payloads move with grouped
// ncclSend / ncclRecv of exact sizes on a side stream while the context computes its next unit.
//
// NCCL is resolved at run time: a process that already loaded libnccl.so.2 (torch) shares that copy, others open the system one.
// =====================================================================================
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "cvx_internal.h"
#include "cvx_kernels.cuh"

namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;
};
NcclApi* nccl_api() {
    static NcclApi api;
    static bool tried = false;
    if (tried) return api.handle ? &api : nullptr;
    tried = true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        api.handle = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);   // the copy the process already has (torch's)
        if (api.handle) break;
    }
    for (const char* n : names) {
        if (api.handle) break;
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!api.handle) { api.err = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "?"); return nullptr; }
#define CVX_SYM(field, name)                                                            \
    *reinterpret_cast<void**>(&api.field) = dlsym(api.handle, name);                    \
    if (!api.field) { api.err = std::string("libnccl lacks ") + name; api.handle = nullptr; return nullptr; }
    CVX_SYM(GetUniqueId, "ncclGetUniqueId") CVX_SYM(CommInitRank, "ncclCommInitRank") CVX_SYM(CommDestroy, "ncclCommDestroy")
    CVX_SYM(AllGather, "ncclAllGather") CVX_SYM(Send, "ncclSend") CVX_SYM(Recv, "ncclRecv") CVX_SYM(GroupStart, "ncclGroupStart")
    CVX_SYM(GroupEnd, "ncclGroupEnd") CVX_SYM(GetErrorString, "ncclGetErrorString")
#undef CVX_SYM
    return &api;
}

constexpr int META = 8;          // int64 per rank and round: {P, view1, view2, h1_begin, h1_end, unit_id, V, S = supporting edges of all rows}
constexpr int SLOTS = 2;         // staging buffers in flight

inline size_t align16(size_t x) { return (x + 15) & ~size_t(15); }
// Byte offsets of the fields of a packed unit with P rows and S supports.  The support indices travel as a CSR list (row r owns
// entries [sup_off[r], sup_off[r+1]), each {validation view slot, edge index}, ascending slot): a row is supported by ~10 of ~100
// views, so the dense P x V matrix of the reference's paired_edge (EdgeSketch_Core.cpp:510-526) would be ~4x the bytes.
struct Layout { size_t h1, h2, cnt, nan, gw, tw, gc, h2c, soff, sent, total; };
Layout layout_of(int64_t P, int64_t S) {
    Layout l;
    size_t o = 0;
    l.h1 = o; o = align16(o + size_t(P) * 4);
    l.h2 = o; o = align16(o + size_t(P) * 4);
    l.cnt = o; o = align16(o + size_t(P) * 4);
    l.nan = o; o = align16(o + size_t(P) * 4);
    l.gw = o; o = align16(o + size_t(P) * 24);
    l.tw = o; o = align16(o + size_t(P) * 24);
    l.gc = o; o = align16(o + size_t(P) * 24);
    l.h2c = o; o = align16(o + size_t(P) * 16);
    l.soff = o; o = align16(o + size_t(P + 1) * 8);
    l.sent = o; o = align16(o + size_t(S) * 8);
    l.total = o;
    return l;
}
// one warp per accepted row: the non-empty entries of its dense support row, in view order
__global__ void __launch_bounds__(256) k_pack_supports(const int32_t* __restrict__ sup, int64_t P, int V, const int64_t* __restrict__ off,
                                                       int2* __restrict__ ent) {
    const int lane = threadIdx.x & 31;
    const int64_t p = blockIdx.x * int64_t(blockDim.x >> 5) + (threadIdx.x >> 5);
    if (p >= P) return;
    int64_t o = off[p];
    for (int v0 = 0; v0 < V; v0 += 32) {
        const int v = v0 + lane;
        const int e = v < V ? sup[p * V + v] : -2;
        const unsigned m = __ballot_sync(0xffffffffu, e != -2);
        if (e != -2) ent[o + __popc(m & ((1u << lane) - 1u))] = make_int2(v, e);
        o += __popc(m);
    }
}
__global__ void k_write_meta(int64_t* meta, int64_t P, int v1, int v2, int rb, int re, int64_t unit, int V, const int64_t* S) {
    meta[0] = P; meta[1] = v1; meta[2] = v2; meta[3] = rb; meta[4] = re; meta[5] = unit; meta[6] = V; meta[7] = (P > 0 && S) ? *S : 0;
}

struct Round {
    int slot = 0;
    bool posted = false;                 // payload transfers enqueued
    cudaEvent_t ev_meta = nullptr;       // meta of all ranks is in h_meta
    int64_t* h_meta = nullptr;           // pinned, world x META
    std::vector<char*> base;             // rank 0: arena address of every rank's block
};

}  // namespace

struct CvxCommState {
    ncclComm_t comm = nullptr;
    bool own_comm = false;
    int world = 1, rank = 0, device = 0;
    cudaStream_t cs = nullptr;           // the side stream every NCCL call of this context is ordered on
    // staging (send side)
    char* stage[SLOTS] = {nullptr, nullptr};
    size_t stage_cap[SLOTS] = {0, 0};
    size_t stage_bytes[SLOTS] = {0, 0};
    cudaEvent_t ev_packed[SLOTS] = {nullptr, nullptr}, ev_sent[SLOTS] = {nullptr, nullptr};
    bool slot_busy[SLOTS] = {false, false};
    int64_t* d_meta_send = nullptr;      // META
    int64_t* d_meta_all = nullptr;       // world x META per slot
    int64_t* h_meta_send = nullptr;      // pinned, SLOTS x META
    std::vector<Round> rounds;          // pool: events and pinned records are reused across cvx_gather_reset
    size_t n_rounds = 0;                // rounds posted since the last reset
    // receive arena (rank 0): chunks are never moved while transfers are in flight
    struct Chunk { char* p; size_t cap, used; };
    std::vector<Chunk> arena;
    int64_t bytes_sent = 0, bytes_recv = 0;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
    bool timing_open = false;
    // claims all-reduce scratch
    uint32_t* d_claim_all = nullptr;
    size_t claim_all_words = 0;
};

namespace {

#define CUC(call)                                                                                                        \
    do {                                                                                                                 \
        cudaError_t _e = (call);                                                                                         \
        if (_e != cudaSuccess) return cvx_internal_fail(ctx, CVX_ERR_CUDA, (std::string(#call) + ": " + cudaGetErrorString(_e)).c_str()); \
    } while (0)
#define NCC(call)                                                                                                        \
    do {                                                                                                                 \
        ncclResult_t _r = (call);                                                                                        \
        if (_r != ncclSuccess) return cvx_internal_fail(ctx, CVX_ERR_CUDA, (std::string(#call) + ": " + api->GetErrorString(_r)).c_str()); \
    } while (0)

CvxCommState* state_of(cvx_ctx* ctx) { return ctx ? *cvx_internal_comm_slot(ctx) : nullptr; }

int make_state(cvx_ctx* ctx, ncclComm_t comm, bool own, int world, int rank) {
    CvxCommState* s = new CvxCommState();
    s->comm = comm; s->own_comm = own; s->world = world; s->rank = rank;
    cudaStream_t main_stream;
    cvx_internal_device(ctx, &s->device, &main_stream);
    CUC(cudaSetDevice(s->device));
    CUC(cudaStreamCreateWithFlags(&s->cs, cudaStreamNonBlocking));
    for (int k = 0; k < SLOTS; k++) {
        CUC(cudaEventCreateWithFlags(&s->ev_packed[k], cudaEventDisableTiming));
        CUC(cudaEventCreateWithFlags(&s->ev_sent[k], cudaEventDisableTiming));
    }
    CUC(cudaEventCreate(&s->ev_t0)); CUC(cudaEventCreate(&s->ev_t1));
    CUC(cudaMalloc(&s->d_meta_send, SLOTS * META * 8));
    CUC(cudaMalloc(&s->d_meta_all, size_t(SLOTS) * world * META * 8));
    CUC(cudaMallocHost(reinterpret_cast<void**>(&s->h_meta_send), SLOTS * META * 8));
    *cvx_internal_comm_slot(ctx) = s;
    return CVX_OK;
}

// rank 0: room for `bytes` in the receive arena (a new chunk when the current ones are full; nothing in flight ever moves)
char* arena_take(cvx_ctx* ctx, CvxCommState* s, size_t bytes) {
    bytes = align16(std::max<size_t>(bytes, 16));
    for (auto& c : s->arena)
        if (c.cap - c.used >= bytes) { char* p = c.p + c.used; c.used += bytes; return p; }
    size_t cap = std::max<size_t>(bytes, size_t(64) << 20);
    if (!s->arena.empty()) cap = std::max(cap, s->arena.back().cap * 2);
    char* p = nullptr;
    if (cudaMalloc(&p, cap) != cudaSuccess) { cvx_internal_fail(ctx, CVX_ERR_NOMEM, "cvx_gather: receive arena allocation failed"); return nullptr; }
    s->arena.push_back({p, cap, bytes});
    return p;
}

// enqueue the payload transfers of round r (its meta must have arrived)
int post_round(cvx_ctx* ctx, CvxCommState* s, NcclApi* api, size_t r) {
    Round& R = s->rounds[r];
    if (R.posted) return CVX_OK;
    CUC(cudaEventSynchronize(R.ev_meta));
    const int k = R.slot;
    const int64_t* m = R.h_meta;
    if (s->rank == 0) {
        R.base.assign(size_t(s->world), nullptr);
        for (int q = 0; q < s->world; q++) {
            const Layout l = layout_of(m[q * META], m[q * META + 7]);
            if (m[q * META] > 0) { R.base[size_t(q)] = arena_take(ctx, s, l.total); if (!R.base[size_t(q)]) return CVX_ERR_NOMEM; }
        }
        NCC(api->GroupStart());
        for (int q = 1; q < s->world; q++)
            if (m[q * META] > 0) {
                const size_t b = layout_of(m[q * META], m[q * META + 7]).total;
                NCC(api->Recv(R.base[size_t(q)], b, ncclChar, q, s->comm, s->cs));
                s->bytes_recv += int64_t(b);
            }
        NCC(api->GroupEnd());
        if (m[0] > 0) CUC(cudaMemcpyAsync(R.base[0], s->stage[k], layout_of(m[0], m[7]).total, cudaMemcpyDeviceToDevice, s->cs));
    } else if (m[s->rank * META] > 0) {
        const size_t b = layout_of(m[s->rank * META], m[s->rank * META + 7]).total;   // exact: the support count came back with the meta
        NCC(api->GroupStart());
        NCC(api->Send(s->stage[k], b, ncclChar, 0, s->comm, s->cs));
        NCC(api->GroupEnd());
        s->bytes_sent += int64_t(b);
    }
    CUC(cudaEventRecord(s->ev_sent[k], s->cs));
    R.posted = true;
    return CVX_OK;
}

}  // namespace

void cvx_comm_release(CvxCommState* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->cs) cudaStreamSynchronize(s->cs);
    if (s->own_comm && s->comm) { NcclApi* api = nccl_api(); if (api) api->CommDestroy(s->comm); }
    for (int k = 0; k < SLOTS; k++) {
        if (s->stage[k]) cudaFree(s->stage[k]);
        if (s->ev_packed[k]) cudaEventDestroy(s->ev_packed[k]);
        if (s->ev_sent[k]) cudaEventDestroy(s->ev_sent[k]);
    }
    for (auto& r : s->rounds) { if (r.ev_meta) cudaEventDestroy(r.ev_meta); if (r.h_meta) cudaFreeHost(r.h_meta); }
    for (auto& c : s->arena) cudaFree(c.p);
    if (s->d_meta_send) cudaFree(s->d_meta_send);
    if (s->d_meta_all) cudaFree(s->d_meta_all);
    if (s->h_meta_send) cudaFreeHost(s->h_meta_send);
    if (s->d_claim_all) cudaFree(s->d_claim_all);
    if (s->ev_t0) cudaEventDestroy(s->ev_t0);
    if (s->ev_t1) cudaEventDestroy(s->ev_t1);
    if (s->cs) cudaStreamDestroy(s->cs);
    delete s;
}

extern "C" {

int cvx_comm_unique_id(void* id128) {
    NcclApi* api = nccl_api();
    if (!api || !id128) return CVX_ERR_INVALID;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (api->GetUniqueId(&id) != ncclSuccess) return CVX_ERR_CUDA;
    std::memcpy(id128, &id, 128);
    return CVX_OK;
}

int cvx_comm_init(cvx_ctx* ctx, int world, int rank, const void* id128) {
    if (!ctx || !id128 || world < 1 || rank < 0 || rank >= world) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_comm_init: bad arguments");
    if (state_of(ctx)) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_comm_init: communicator already set");
    NcclApi* api = nccl_api();
    if (!api) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_comm_init: NCCL unavailable (libnccl.so.2 not found)");
    int device; cudaStream_t st;
    cvx_internal_device(ctx, &device, &st);
    CUC(cudaSetDevice(device));
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    ncclComm_t comm;
    NCC(api->CommInitRank(&comm, world, id, rank));
    return make_state(ctx, comm, true, world, rank);
}

int cvx_comm_attach(cvx_ctx* ctx, void* nccl_comm, int world, int rank) {
    if (!ctx || !nccl_comm || world < 1 || rank < 0 || rank >= world) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_comm_attach: bad arguments");
    if (state_of(ctx)) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_comm_attach: communicator already set");
    if (!nccl_api()) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_comm_attach: NCCL unavailable");
    return make_state(ctx, static_cast<ncclComm_t>(nccl_comm), false, world, rank);
}

int cvx_comm_destroy(cvx_ctx* ctx) {
    if (!ctx) return CVX_ERR_INVALID;
    CvxCommState** slot = cvx_internal_comm_slot(ctx);
    cvx_comm_release(*slot);
    *slot = nullptr;
    return CVX_OK;
}

int cvx_gather_reset(cvx_ctx* ctx) {
    CvxCommState* s = state_of(ctx);
    if (!s) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_gather_reset: no communicator");
    CUC(cudaSetDevice(s->device));
    CUC(cudaStreamSynchronize(s->cs));
    for (auto& c : s->arena) c.used = 0;
    for (auto& r : s->rounds) r.posted = false;
    s->n_rounds = 0;
    s->bytes_sent = s->bytes_recv = 0;
    s->timing_open = false;
    for (int k = 0; k < SLOTS; k++) s->slot_busy[k] = false;
    return CVX_OK;
}

int cvx_gather_post(cvx_ctx* ctx, int64_t unit_id, int empty) {
    CvxCommState* s = state_of(ctx);
    NcclApi* api = nccl_api();
    if (!s || !api) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_gather_post: no communicator");
    CUC(cudaSetDevice(s->device));
    const size_t r = s->n_rounds;
    // payloads of the earlier rounds first: their meta arrived long ago (a whole unit of compute later), and the NCCL call order
    // p2p(r-1), all-gather(r) is then the same on every rank
    for (size_t q = 0; q < r; q++) { const int rc = post_round(ctx, s, api, q); if (rc) return rc; }
    const int k = int(r % SLOTS);
    if (s->slot_busy[k]) CUC(cudaEventSynchronize(s->ev_sent[k]));   // the staging buffer of round r - SLOTS is free again
    CvxResultsView v;
    std::memset(&v, 0, sizeof(v));
    cudaStream_t main_stream;
    int dev;
    cvx_internal_device(ctx, &dev, &main_stream);
    if (!empty) { const int rc = cvx_internal_results(ctx, &v); if (rc) return rc; }
    const Layout l = layout_of(v.P, v.P * int64_t(v.V));          // capacity: every support slot taken
    int64_t* d_S = nullptr;
    if (v.P > 0) {
        if (l.total > s->stage_cap[k]) {
            if (s->stage[k]) CUC(cudaFree(s->stage[k]));
            s->stage[k] = nullptr;
            const size_t cap = l.total + l.total / 4 + 4096;
            CUC(cudaMalloc(&s->stage[k], cap));
            s->stage_cap[k] = cap;
        }
        char* b = s->stage[k];
        const size_t P = size_t(v.P);
        CUC(cudaMemcpyAsync(b + l.h1, v.h1, P * 4, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.h2, v.h2rep, P * 4, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.cnt, v.count, P * 4, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.nan, v.nan, P * 4, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.gw, v.gw, P * 24, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.tw, v.tw, P * 24, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.gc, v.gc, P * 24, cudaMemcpyDeviceToDevice, main_stream));
        CUC(cudaMemcpyAsync(b + l.h2c, v.h2c, P * 16, cudaMemcpyDeviceToDevice, main_stream));
        // supports as a CSR list: offsets = exclusive scan of the per-row support counts (the vote of EdgeSketch_Core.cpp:414-427)
        int64_t* off = reinterpret_cast<int64_t*>(b + l.soff);
        cvx::launch_scan_i32_i64(v.count, off, int(v.P), main_stream);
        k_pack_supports<<<unsigned((v.P + 7) / 8), 256, 0, main_stream>>>(v.sup, v.P, v.V, off, reinterpret_cast<int2*>(b + l.sent));
        cvx_internal_count_launch(ctx, 2);
        d_S = off + v.P;
    }
    // the 64-byte record of this rank (the support count is only known on the device), all-gathered
    k_write_meta<<<1, 1, 0, main_stream>>>(s->d_meta_send + k * META, v.P, v.view1, v.view2, v.run_begin, v.run_end, unit_id, v.V, d_S);
    CUC(cudaGetLastError());
    CUC(cudaEventRecord(s->ev_packed[k], main_stream));
    CUC(cudaStreamWaitEvent(s->cs, s->ev_packed[k], 0));
    if (!s->timing_open) { CUC(cudaEventRecord(s->ev_t0, s->cs)); s->timing_open = true; }
    if (r == s->rounds.size()) {
        Round N;
        CUC(cudaEventCreateWithFlags(&N.ev_meta, cudaEventDisableTiming));
        CUC(cudaMallocHost(reinterpret_cast<void**>(&N.h_meta), size_t(s->world) * META * 8));
        s->rounds.push_back(N);
    }
    Round& R = s->rounds[r];
    R.slot = k; R.posted = false;
    int64_t* d_all = s->d_meta_all + size_t(k) * s->world * META;
    NCC(api->AllGather(s->d_meta_send + k * META, d_all, META, ncclInt64, s->comm, s->cs));
    CUC(cudaMemcpyAsync(R.h_meta, d_all, size_t(s->world) * META * 8, cudaMemcpyDeviceToHost, s->cs));
    CUC(cudaEventRecord(R.ev_meta, s->cs));
    s->slot_busy[k] = true;
    s->n_rounds = r + 1;
    return CVX_OK;
}

int cvx_gather_wait(cvx_ctx* ctx, int join_stream, int host_sync) {
    CvxCommState* s = state_of(ctx);
    NcclApi* api = nccl_api();
    if (!s || !api) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_gather_wait: no communicator");
    CUC(cudaSetDevice(s->device));
    for (size_t q = 0; q < s->n_rounds; q++) { const int rc = post_round(ctx, s, api, q); if (rc) return rc; }
    CUC(cudaEventRecord(s->ev_t1, s->cs));
    if (join_stream) {
        cudaStream_t main_stream; int dev;
        cvx_internal_device(ctx, &dev, &main_stream);
        CUC(cudaStreamWaitEvent(main_stream, s->ev_t1, 0));
    }
    if (host_sync) CUC(cudaStreamSynchronize(s->cs));
    return CVX_OK;
}

int cvx_gather_get(cvx_ctx* ctx, int round, int rank, cvx_gathered_unit* out) {
    CvxCommState* s = state_of(ctx);
    if (!s || !out || s->rank != 0 || round < 0 || size_t(round) >= s->n_rounds || rank < 0 || rank >= s->world)
        return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_gather_get: bad arguments (rank 0 only, after cvx_gather_wait)");
    const Round& R = s->rounds[size_t(round)];
    if (!R.posted) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_gather_get: round not complete (call cvx_gather_wait)");
    std::memset(out, 0, sizeof(*out));
    std::memcpy(out->meta, R.h_meta + rank * META, META * 8);
    const int64_t P = out->meta[0];
    if (P <= 0) return CVX_OK;
    const Layout l = layout_of(P, out->meta[7]);
    const char* b = R.base[size_t(rank)];
    out->pair_h1 = reinterpret_cast<const int32_t*>(b + l.h1); out->pair_h2rep = reinterpret_cast<const int32_t*>(b + l.h2);
    out->pair_count = reinterpret_cast<const int32_t*>(b + l.cnt); out->nan_flag = reinterpret_cast<const int32_t*>(b + l.nan);
    out->support_off = reinterpret_cast<const int64_t*>(b + l.soff); out->support_entries = reinterpret_cast<const int32_t*>(b + l.sent);
    out->gamma_world = reinterpret_cast<const double*>(b + l.gw); out->tangent_world = reinterpret_cast<const double*>(b + l.tw);
    out->gamma_cam = reinterpret_cast<const double*>(b + l.gc); out->h2_corrected_xy = reinterpret_cast<const double*>(b + l.h2c);
    return CVX_OK;
}

static int gather_fetch_impl(cvx_ctx* ctx, int round, int rank, int32_t* h1, int32_t* h2rep, int32_t* count, int32_t* nan_flag, int64_t* support_off,
                             int32_t* support_entries, double* gamma_world, double* tangent_world, double* gamma_cam, double* h2c, bool wait) {
    cvx_gathered_unit u;
    const int rc = cvx_gather_get(ctx, round, rank, &u);
    if (rc) return rc;
    CvxCommState* s = state_of(ctx);
    const size_t P = size_t(u.meta[0]), S = size_t(u.meta[7]);
    if (P == 0) return CVX_OK;
    CUC(cudaSetDevice(s->device));
#define CVX_FETCH(dst, src, bytes) if (dst) CUC(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s->cs))
    CVX_FETCH(h1, u.pair_h1, P * 4); CVX_FETCH(h2rep, u.pair_h2rep, P * 4); CVX_FETCH(count, u.pair_count, P * 4); CVX_FETCH(nan_flag, u.nan_flag, P * 4);
    CVX_FETCH(support_off, u.support_off, (P + 1) * 8); CVX_FETCH(support_entries, u.support_entries, S * 8);
    CVX_FETCH(gamma_world, u.gamma_world, P * 24); CVX_FETCH(tangent_world, u.tangent_world, P * 24); CVX_FETCH(gamma_cam, u.gamma_cam, P * 24);
    CVX_FETCH(h2c, u.h2_corrected_xy, P * 16);
#undef CVX_FETCH
    if (wait) CUC(cudaStreamSynchronize(s->cs));
    return CVX_OK;
}
int cvx_gather_fetch(cvx_ctx* ctx, int round, int rank, int32_t* h1, int32_t* h2rep, int32_t* count, int32_t* nan_flag, int64_t* support_off,
                     int32_t* support_entries, double* gamma_world, double* tangent_world, double* gamma_cam, double* h2_corrected_xy) {
    return gather_fetch_impl(ctx, round, rank, h1, h2rep, count, nan_flag, support_off, support_entries, gamma_world, tangent_world, gamma_cam, h2_corrected_xy, true);
}
int cvx_gather_fetch_async(cvx_ctx* ctx, int round, int rank, int32_t* h1, int32_t* h2rep, int32_t* count, int32_t* nan_flag, int64_t* support_off,
                           int32_t* support_entries, double* gamma_world, double* tangent_world, double* gamma_cam, double* h2_corrected_xy) {
    return gather_fetch_impl(ctx, round, rank, h1, h2rep, count, nan_flag, support_off, support_entries, gamma_world, tangent_world, gamma_cam, h2_corrected_xy, false);
}

int cvx_gather_stats(cvx_ctx* ctx, int64_t* bytes_sent, int64_t* bytes_received, float* comm_ms) {
    CvxCommState* s = state_of(ctx);
    if (!s) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_gather_stats: no communicator");
    if (bytes_sent) *bytes_sent = s->bytes_sent;
    if (bytes_received) *bytes_received = s->bytes_recv;
    if (comm_ms) {
        *comm_ms = 0.f;
        if (s->timing_open) { CUC(cudaSetDevice(s->device)); CUC(cudaEventSynchronize(s->ev_t1)); cudaEventElapsedTime(comm_ms, s->ev_t0, s->ev_t1); }
    }
    return CVX_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------- claims: bitwise OR over the ranks
namespace {
__global__ void k_or_words(uint32_t* __restrict__ dst, const uint32_t* __restrict__ all, size_t n, int world) {
    const size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    uint32_t w = 0;
    for (int r = 0; r < world; r++) w |= all[size_t(r) * n + i];
    dst[i] = w;
}
}  // namespace

extern "C" int cvx_claims_allreduce(cvx_ctx* ctx) {
    CvxCommState* s = state_of(ctx);
    NcclApi* api = nccl_api();
    if (!s || !api) return cvx_internal_fail(ctx, CVX_ERR_INVALID, "cvx_claims_allreduce: no communicator");
    CUC(cudaSetDevice(s->device));
    uint32_t* words = nullptr;
    size_t n = 0;
    const int rc = cvx_internal_claim_words(ctx, &words, &n);
    if (rc) return rc;
    if (s->world == 1 || n == 0) return CVX_OK;
    if (s->claim_all_words < n * size_t(s->world)) {
        if (s->d_claim_all) CUC(cudaFree(s->d_claim_all));
        s->d_claim_all = nullptr;
        CUC(cudaMalloc(&s->d_claim_all, n * size_t(s->world) * 4));
        s->claim_all_words = n * size_t(s->world);
    }
    cudaStream_t main_stream; int dev;
    cvx_internal_device(ctx, &dev, &main_stream);
    // on the context stream: the claim kernels that precede it and the count kernels that follow are ordered there
    NCC(api->AllGather(words, s->d_claim_all, n, ncclUint32, s->comm, main_stream));
    k_or_words<<<unsigned((n + 255) / 256), 256, 0, main_stream>>>(words, s->d_claim_all, n, s->world);
    cvx_internal_count_launch(ctx, 1);
    CUC(cudaGetLastError());
    return CVX_OK;
}
