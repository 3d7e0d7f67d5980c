// comm.cu -- the communicator of the C-ABI (include/ntxent_b200.h, "Communicator"): what a rank
// owns between its shard and its gradient in the row-partitioned step, and the step itself as
// one call.  The reference has no counterpart: it names NCCL and MPI as build options
// (CMakeLists.txt:13-14,41-47,115-121) and calls neither.
//
// Two ways to move the data, both behind ntxent_b200_dist_forward_backward:
//   PUSH  the kernels store U, l and the partial loss straight into every rank's buffers and order
//         themselves through flags (ntxent_b200_normalize_push / _forward_rows_push /
//         _backward_rows_dz); the buffers are plain cudaMalloc memory that every peer maps -- peer
//         access inside one process, CUDA IPC handles across processes.  No collective launches.
//   NCCL  north_star's wording: ncclAllGather of U, then either ncclAllGather of l (symmetric
//         backward) or ncclReduceScatter of the column term (contract backward) on a second stream
//         while the row term is computed on the first.
// NCCL is loaded with dlopen at communicator creation (the library keeps no link-time dependency
// on it); without it only single-process PUSH communicators can be made.
#include "../../include/ntxent_b200.h"

#include <dlfcn.h>
#include <nccl.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"

namespace ntxent_b200 {
// ntxent_b200_backward_rows_dz with a caller-owned ticket array for the in-kernel epilogue (capi.cu)
int backward_rows_dz_tickets(const void* u_all, const float* lse_all, int64_t M, int64_t D, int64_t row_begin,
                             int64_t row_count, float tau, int dtype, unsigned flags, void* workspace,
                             size_t workspace_bytes, const void* z_rows, const float* inv_norm_rows,
                             const float* grad_out, void* dz_out, const ntxent_peer_wait* lse_ready,
                             const float* loss_parts, int loss_parts_count, float* loss_out, unsigned* tickets,
                             cudaStream_t stream);
}  // namespace ntxent_b200


namespace ntxent_b200 {
namespace {

// ------------------------------------------------------------------- NCCL, late-bound ---
struct Nccl {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                                  cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

// A process that already has a NCCL loaded (PyTorch brings its own) gets that copy: same soname.
const Nccl& nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {getenv("NTXENT_NCCL_LIBRARY"), "libnccl.so.2", "libnccl.so"};
        for (const char* name : names) {
            if (!name || !name[0]) continue;
            n.handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (n.handle) break;
        }
        if (!n.handle) return;
#define NTX_SYM(field, sym) n.field = reinterpret_cast<decltype(n.field)>(dlsym(n.handle, sym))
        NTX_SYM(GetUniqueId, "ncclGetUniqueId");
        NTX_SYM(CommInitRank, "ncclCommInitRank");
        NTX_SYM(CommInitAll, "ncclCommInitAll");
        NTX_SYM(CommDestroy, "ncclCommDestroy");
        NTX_SYM(AllGather, "ncclAllGather");
        NTX_SYM(ReduceScatter, "ncclReduceScatter");
        NTX_SYM(AllReduce, "ncclAllReduce");
        NTX_SYM(GetErrorString, "ncclGetErrorString");
#undef NTX_SYM
        n.ok = n.GetUniqueId && n.CommInitRank && n.CommInitAll && n.CommDestroy && n.AllGather &&
               n.ReduceScatter && n.AllReduce;
    });
    return n;
}

inline bool debug_on() {
    static const bool on = [] { const char* v = getenv("NTXENT_DEBUG"); return v && v[0] == '1'; }();
    return on;
}
#define NTX_CUDA(expr)                                                                             \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            if (debug_on())                                                                        \
                fprintf(stderr, "[ntxent_b200] %s:%d: %s -> %s\n", __FILE__, __LINE__, #expr,      \
                        cudaGetErrorString(e__));                                                  \
            return (int)e__;                                                                       \
        }                                                                                          \
    } while (0)
#define NTX_NCCL(expr)                                                                             \
    do {                                                                                           \
        ncclResult_t r__ = (expr);                                                                 \
        if (r__ != ncclSuccess) {                                                                  \
            if (debug_on())                                                                        \
                fprintf(stderr, "[ntxent_b200] %s:%d: %s -> NCCL error %d (%s)\n", __FILE__, __LINE__, #expr, \
                        (int)r__, nccl().GetErrorString ? nccl().GetErrorString(r__) : "?");       \
            return NTXENT_ERR_COLLECTIVE;                                                          \
        }                                                                                          \
    } while (0)
#define NTX_RC(expr)                    \
    do {                                \
        int rc__ = (expr);              \
        if (rc__ != NTXENT_OK) return rc__; \
    } while (0)

constexpr size_t kAl = 256;
inline size_t al(size_t n) { return (n + kAl - 1) / kAl * kAl; }

// sync block of a buffer set (256 bytes): flag_u[16] | flag_lse[16] | epoch_u | . | epoch_lse | . | status
constexpr size_t kOffFlagU = 0, kOffFlagLse = 64, kOffEpochU = 128, kOffEpochLse = 136, kOffStatus = 144;

}  // namespace
}  // namespace ntxent_b200

using namespace ntxent_b200;

struct ntxent_comm {
    int dev = 0, rank = 0, nranks = 1;
    int64_t M = 0, D = 0, rows = 0, row_begin = 0;
    int dtype = 0, op_dtype = 0;
    unsigned flags = 0;
    size_t item = 2, u_bytes = 0, lse_bytes = 0, set_bytes = 0;
    char* base = nullptr;                       // this rank's two buffer sets
    char* peer[NTXENT_MAX_PEERS] = {};          // every rank's `base` as mapped here (own included)
    bool ipc_opened[NTXENT_MAX_PEERS] = {};
    void* ws = nullptr;
    size_t ws_bytes = 0;
    float* inv_norm = nullptr;
    unsigned* tickets = nullptr;                // 1 + rows/128 words, zero between calls (in-kernel finalize)
    float* loss_part = nullptr;                 // [2]: partial, summed (NCCL paths)
    float* du_rows = nullptr;                   // contract backward: row term [rows, D]
    float* du_cols = nullptr;                   //                    column term [M, D]
    float* du_col_local = nullptr;              //                    its reduce-scattered share [rows, D]
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    ncclComm_t nccl_comm = nullptr;
    unsigned step = 0;
};

namespace ntxent_b200 {
namespace {

// loss = scale * sum of the ranks' partial losses, in rank order (the same bits on every rank)
__global__ void __launch_bounds__(32) loss_sum_kernel(const float* parts, int n, float scale, float* out) {
    pdl_trigger();
    pdl_wait();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int i = 0; i < n; ++i) s += parts[i];
        *out = s * scale;
    }
}
// a += b (the reduce-scattered column term onto the row term)
__global__ void __launch_bounds__(256) add_inplace_kernel(float* __restrict__ a, const float* __restrict__ b,
                                                          int64_t n4) {
    pdl_trigger();
    pdl_wait();
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n4) {
        float4 x = reinterpret_cast<float4*>(a)[i];
        const float4 y = reinterpret_cast<const float4*>(b)[i];
        x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
        reinterpret_cast<float4*>(a)[i] = x;
    }
}

int check_problem(int nranks, int64_t M, int64_t D, int dtype, unsigned flags) {
    if (nranks < 1 || nranks > NTXENT_MAX_PEERS) return NTXENT_ERR_INVALID_ARGUMENT;
    if (flags & NTXENT_FLAG_CROSS_VIEW) return NTXENT_ERR_INVALID_ARGUMENT;  // single-device entries only
    if (dtype != NTXENT_F32 && dtype != NTXENT_BF16) return NTXENT_ERR_INVALID_ARGUMENT;
    if (M < 2 || (M & 1) || M % nranks) return NTXENT_ERR_INVALID_ARGUMENT;
    if (D < 8 || (D % 8) != 0 || D > 1024 || M > ((int64_t)1 << 24)) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    if (ntxent_b200_rows_workspace_bytes(M, D, M / nranks, dtype, 0) == 0) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    return NTXENT_OK;
}

void free_comm(ntxent_comm* c) {
    if (!c) return;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(c->dev);
    cudaDeviceSynchronize();
    for (int r = 0; r < c->nranks; ++r)
        if (c->ipc_opened[r] && c->peer[r]) cudaIpcCloseMemHandle(c->peer[r]);
    if (c->nccl_comm && nccl().ok) nccl().CommDestroy(c->nccl_comm);
    if (c->side) cudaStreamDestroy(c->side);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    for (void* p : {(void*)c->base, c->ws, (void*)c->inv_norm, (void*)c->tickets, (void*)c->loss_part, (void*)c->du_rows,
                    (void*)c->du_cols, (void*)c->du_col_local})
        if (p) cudaFree(p);
    (void)cudaGetLastError();
    cudaSetDevice(prev);
    delete c;
}

// Everything a rank owns, on the current device (= c->dev).  The buffer sets start zeroed (flags,
// call counters, status words).
int alloc_comm(ntxent_comm* c, int rank, int nranks, int64_t M, int64_t D, int dtype, unsigned flags) {
    c->rank = rank; c->nranks = nranks; c->M = M; c->D = D; c->dtype = dtype; c->flags = flags;
    c->rows = M / nranks;
    c->row_begin = (int64_t)rank * c->rows;
    c->item = (size_t)ntxent_b200_operand_size(dtype, flags);
    c->op_dtype = c->item == 2 ? NTXENT_BF16 : NTXENT_F32;
    c->u_bytes = al((size_t)M * D * c->item);
    c->lse_bytes = al((size_t)M * 4);
    c->set_bytes = c->u_bytes + c->lse_bytes + 2 * kAl;
    NTX_CUDA(cudaMalloc((void**)&c->base, 2 * c->set_bytes));
    NTX_CUDA(cudaMemset(c->base, 0, 2 * c->set_bytes));
    c->ws_bytes = ntxent_b200_rows_workspace_bytes(M, D, c->rows, dtype, flags);
    NTX_CUDA(cudaMalloc(&c->ws, c->ws_bytes));
    NTX_CUDA(cudaMalloc((void**)&c->inv_norm, al(sizeof(float) * c->rows)));
    // tickets: [0] the loss, [1 .. T] the forward's row tiles, [1 + T] the normalise kernel's signal,
    // [2 + T .. 2 + 2T) the backward's row tiles (in-kernel epilogue); every user leaves its words zero
    const size_t tick_bytes = al(sizeof(unsigned) * (2 + 2 * (size_t)((c->rows + 127) / 128)));
    NTX_CUDA(cudaMalloc((void**)&c->tickets, tick_bytes));
    NTX_CUDA(cudaMemset(c->tickets, 0, tick_bytes));
    NTX_CUDA(cudaMalloc((void**)&c->loss_part, kAl));
    NTX_CUDA(cudaMalloc((void**)&c->du_rows, al(sizeof(float) * c->rows * D)));
    NTX_CUDA(cudaMalloc((void**)&c->du_col_local, al(sizeof(float) * c->rows * D)));
    NTX_CUDA(cudaMalloc((void**)&c->du_cols, al(sizeof(float) * M * D)));
    NTX_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
    NTX_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    NTX_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    NTX_CUDA(cudaDeviceSynchronize());
    c->peer[rank] = c->base;
    return NTXENT_OK;
}

}  // namespace
}  // namespace ntxent_b200

extern "C" {

int ntxent_b200_comm_init_all(int ndev, const int* devices, int64_t M, int64_t D, int dtype, unsigned flags,
                              ntxent_comm_t** comms_out) {
    if (!devices || !comms_out) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_problem(ndev, M, D, dtype, flags);
    if (rc != NTXENT_OK) return rc;
    int prev = 0;
    NTX_CUDA(cudaGetDevice(&prev));
    std::vector<ntxent_comm*> cs(ndev, nullptr);
    auto fail = [&](int code) {
        for (auto* c : cs) free_comm(c);
        cudaSetDevice(prev);
        return code;
    };
    for (int r = 0; r < ndev; ++r) {
        cs[r] = new (std::nothrow) ntxent_comm();
        if (!cs[r]) return fail((int)cudaErrorMemoryAllocation);
        cs[r]->dev = devices[r];
        for (int q = 0; q < r; ++q)  // one rank per GPU: kernels of two ranks on one GPU would wait for
            if (devices[q] == devices[r]) return fail(NTXENT_ERR_INVALID_ARGUMENT);  // each other
        if (cudaSetDevice(devices[r]) != cudaSuccess) return fail(NTXENT_ERR_INVALID_ARGUMENT);
        if (ntxent_b200_device_supported() != 1) return fail(NTXENT_ERR_UNSUPPORTED_DEVICE);
        rc = alloc_comm(cs[r], r, ndev, M, D, dtype, flags);
        if (rc != NTXENT_OK) return fail(rc);
    }
    // one address space: a peer's buffer is reachable as soon as peer access is on
    for (int r = 0; r < ndev; ++r) {
        cudaSetDevice(devices[r]);
        for (int q = 0; q < ndev; ++q) {
            cs[r]->peer[q] = cs[q]->base;
            if (q == r) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devices[r], devices[q]);
            if (!can) return fail(NTXENT_ERR_NO_PEER_ACCESS);
            cudaError_t e = cudaDeviceEnablePeerAccess(devices[q], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail((int)e);
            (void)cudaGetLastError();
        }
    }
    if (ndev > 1 && nccl().ok) {  // optional: the NCCL exchanges
        std::vector<ncclComm_t> nc(ndev, nullptr);
        if (nccl().CommInitAll(nc.data(), ndev, devices) == ncclSuccess)
            for (int r = 0; r < ndev; ++r) cs[r]->nccl_comm = nc[r];
    }
    for (int r = 0; r < ndev; ++r) comms_out[r] = cs[r];
    cudaSetDevice(prev);
    return NTXENT_OK;
}

int ntxent_b200_comm_unique_id(void* id128_out) {
    static_assert(sizeof(ncclUniqueId) == 128, "the id is handed around as 128 opaque bytes");
    if (!id128_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!nccl().ok) return NTXENT_ERR_COLLECTIVE;
    ncclUniqueId id;
    NTX_NCCL(nccl().GetUniqueId(&id));
    memcpy(id128_out, &id, sizeof(id));
    return NTXENT_OK;
}

int ntxent_b200_comm_init_rank(const void* id128, int nranks, int rank, int64_t M, int64_t D, int dtype,
                               unsigned flags, ntxent_comm_t** comm_out) {
    if (!id128 || !comm_out || rank < 0 || rank >= nranks) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_problem(nranks, M, D, dtype, flags);
    if (rc != NTXENT_OK) return rc;
    if (ntxent_b200_device_supported() != 1) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    if (nranks > 1 && !nccl().ok) return NTXENT_ERR_COLLECTIVE;  // the bootstrap runs over NCCL
    ntxent_comm* c = new (std::nothrow) ntxent_comm();
    if (!c) return (int)cudaErrorMemoryAllocation;
    auto fail = [&](int code) { free_comm(c); return code; };
    if (cudaGetDevice(&c->dev) != cudaSuccess) return fail(NTXENT_ERR_UNSUPPORTED_DEVICE);
    rc = alloc_comm(c, rank, nranks, M, D, dtype, flags);
    if (rc != NTXENT_OK) return fail(rc);
    if (nranks > 1) {
        ncclUniqueId id;
        memcpy(&id, id128, sizeof(id));
        if (nccl().CommInitRank(&c->nccl_comm, nranks, id, rank) != ncclSuccess) return fail(NTXENT_ERR_COLLECTIVE);
        // every rank's IPC handle to every rank: an all-gather of 64 bytes each (which is also the
        // barrier behind which all the buffer sets are known to be zeroed)
        cudaIpcMemHandle_t mine, all[NTXENT_MAX_PEERS];
        char* dbuf = nullptr;
        if (cudaIpcGetMemHandle(&mine, c->base) != cudaSuccess) return fail(NTXENT_ERR_NO_PEER_ACCESS);
        if (cudaMalloc((void**)&dbuf, sizeof(mine) * nranks) != cudaSuccess) return fail((int)cudaErrorMemoryAllocation);
        cudaMemcpy(dbuf + sizeof(mine) * rank, &mine, sizeof(mine), cudaMemcpyHostToDevice);
        ncclResult_t nr = nccl().AllGather(dbuf + sizeof(mine) * rank, dbuf, sizeof(mine), ncclChar, c->nccl_comm,
                                           (cudaStream_t)0);
        cudaError_t ce = cudaStreamSynchronize((cudaStream_t)0);
        if (nr == ncclSuccess && ce == cudaSuccess)
            ce = cudaMemcpy(all, dbuf, sizeof(mine) * nranks, cudaMemcpyDeviceToHost);
        cudaFree(dbuf);
        if (nr != ncclSuccess || ce != cudaSuccess) return fail(NTXENT_ERR_COLLECTIVE);
        for (int q = 0; q < nranks; ++q) {
            if (q == rank) continue;
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                (void)cudaGetLastError();
                return fail(NTXENT_ERR_NO_PEER_ACCESS);
            }
            c->peer[q] = (char*)p;
            c->ipc_opened[q] = true;
        }
    }
    *comm_out = c;
    return NTXENT_OK;
}

int ntxent_b200_comm_destroy(ntxent_comm_t* comm) {
    if (!comm) return NTXENT_ERR_INVALID_ARGUMENT;
    free_comm(comm);
    return NTXENT_OK;
}

int ntxent_b200_comm_info(const ntxent_comm_t* comm, int* rank, int* nranks, int64_t* rows_per_rank,
                          int* has_nccl) {
    if (!comm) return NTXENT_ERR_INVALID_ARGUMENT;
    if (rank) *rank = comm->rank;
    if (nranks) *nranks = comm->nranks;
    if (rows_per_rank) *rows_per_rank = comm->rows;
    if (has_nccl) *has_nccl = (comm->nccl_comm != nullptr || comm->nranks == 1) ? 1 : 0;
    return NTXENT_OK;
}

int ntxent_b200_comm_status(ntxent_comm_t* comm) {
    if (!comm) return NTXENT_ERR_INVALID_ARGUMENT;
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(comm->dev);
    unsigned worst = 0;
    for (int k = 0; k < 2; ++k) {
        unsigned v = 0;
        char* sync = comm->base + k * comm->set_bytes + comm->u_bytes + comm->lse_bytes + kAl;
        if (cudaMemcpy(&v, sync + kOffStatus, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) {
            cudaSetDevice(prev);
            return (int)cudaGetLastError();
        }
        if (v) worst = v;
    }
    cudaSetDevice(prev);
    return (int)worst;
}

}  // extern "C"

// ------------------------------------------------------------------------- the step ---
namespace ntxent_b200 {
namespace {

struct SetView {
    char* u_all;
    float* lse_all;
    float* loss_parts;
    ntxent_peer_list u_peers, lse_peers, loss_peers;
    ntxent_peer_signal sig_u, sig_lse;
    ntxent_peer_wait wait_u, wait_lse;
};

SetView view_of(const ntxent_comm* c, int k) {
    SetView v{};
    const size_t o = (size_t)k * c->set_bytes, o_lse = o + c->u_bytes, o_loss = o_lse + c->lse_bytes,
                 o_sync = o_loss + kAl;
    v.u_all = c->base + o;
    v.lse_all = (float*)(c->base + o_lse);
    v.loss_parts = (float*)(c->base + o_loss);
    v.u_peers.count = v.lse_peers.count = v.loss_peers.count = c->nranks;
    v.sig_u.flags.count = v.sig_lse.flags.count = c->nranks;
    for (int r = 0; r < c->nranks; ++r) {
        v.u_peers.ptr[r] = c->peer[r] + o;
        v.lse_peers.ptr[r] = c->peer[r] + o_lse;
        v.loss_peers.ptr[r] = c->peer[r] + o_loss;
        v.sig_u.flags.ptr[r] = c->peer[r] + o_sync + kOffFlagU;
        v.sig_lse.flags.ptr[r] = c->peer[r] + o_sync + kOffFlagLse;
    }
    char* sync = c->base + o_sync;
    v.sig_u.epoch = (uint32_t*)(sync + kOffEpochU);
    v.sig_u.rank = c->rank;
    // The normalise kernel CAN publish the signal itself (its last block, after a system-scope fence
    // per block).  Measured slower than the one-block signal kernel after it (0.415 vs 0.407 ms at
    // C4 on 2 GPUs, profiles/r2_fusion_ab.txt): off unless NTXENT_NORM_TICKET=1.
    static const bool norm_ticket = [] { const char* e = getenv("NTXENT_NORM_TICKET"); return e && e[0] == '1'; }();
    v.sig_u.ticket = norm_ticket ? c->tickets + 1 + (c->rows + 127) / 128 : nullptr;
    v.sig_lse.epoch = (uint32_t*)(sync + kOffEpochLse);
    v.sig_lse.rank = c->rank;
    v.sig_lse.ticket = c->tickets;
    v.wait_u.flags = (const uint32_t*)(sync + kOffFlagU);
    v.wait_u.epoch = (const uint32_t*)(sync + kOffEpochU);
    v.wait_lse.flags = (const uint32_t*)(sync + kOffFlagLse);
    v.wait_lse.epoch = (const uint32_t*)(sync + kOffEpochLse);
    v.wait_u.ranks = v.wait_lse.ranks = c->nranks;
    v.wait_u.rows_per_rank = v.wait_lse.rows_per_rank = c->rows;
    v.wait_u.status = v.wait_lse.status = (uint32_t*)(sync + kOffStatus);
    return v;
}

// PUSH: normalise -> signal -> forward tiles (wait flags) -> finalize (one kernel: positive logits, l,
// loss, signal by its last block) -> backward tiles (wait flags) -> epilogue (+ global loss): six
// launches, no collective.  (Folding the signal into the normalise kernel and the finalize into the
// tile kernel's tail were both built and measured slower: profiles/r2_fusion_ab.txt.)  Two buffer sets alternate so that a fast rank's step t+1 never
// lands in what a slow rank still reads for step t.
int step_push(ntxent_comm* c, const void* z, const float* g, float tau, float* loss_out, void* dz_out,
              cudaStream_t st) {
    const SetView v = view_of(c, (int)(c->step & 1u));
    c->step++;
    const bool multi = c->nranks > 1;
    NTX_RC(ntxent_b200_normalize_push(z, c->rows, c->D, c->dtype, c->flags, &v.u_peers, c->row_begin, c->inv_norm,
                                      multi ? &v.sig_u : nullptr, st));
    NTX_RC(ntxent_b200_forward_rows_push(v.u_all, c->M, c->D, c->row_begin, c->rows, tau, c->op_dtype, c->flags,
                                         c->ws, c->ws_bytes, &v.lse_peers, &v.loss_peers, c->rank,
                                         multi ? &v.wait_u : nullptr, multi ? &v.sig_lse : nullptr, st));
    // (the backward tile kernel waits for every rank's l, hence for every partial loss: its
    // epilogue sums them)
    NTX_RC(backward_rows_dz_tickets(v.u_all, v.lse_all, c->M, c->D, c->row_begin, c->rows, tau, c->dtype,
                                    c->flags, c->ws, c->ws_bytes, z, c->inv_norm, g, dz_out,
                                    multi ? &v.wait_lse : nullptr, v.loss_parts, c->nranks, loss_out,
                                    c->tickets + 2 + (c->rows + 127) / 128, st));
    return NTXENT_OK;
}

// NCCL: the collectives north_star names.  Set 0 only: a collective is its own barrier.
int step_nccl(ntxent_comm* c, const void* z, const float* g, float tau, int bwd_mode, float* loss_out,
              void* dz_out, cudaStream_t st) {
    const SetView v = view_of(c, 0);
    const Nccl& n = nccl();
    const bool multi = c->nranks > 1;
    if (multi && (!n.ok || !c->nccl_comm)) return NTXENT_ERR_COLLECTIVE;
    const ncclDataType_t op_t = c->item == 2 ? ncclBfloat16 : ncclFloat32;
    char* u_mine = v.u_all + (size_t)c->row_begin * c->D * c->item;
    float* lse_mine = v.lse_all + c->row_begin;
    // normalised shard straight into its place in u_all, then an in-place all-gather
    NTX_RC(ntxent_b200_normalize(z, c->rows, c->D, c->dtype, c->flags, u_mine, c->inv_norm, st));
    if (multi) NTX_NCCL(n.AllGather(u_mine, v.u_all, (size_t)c->rows * c->D, op_t, c->nccl_comm, st));
    NTX_RC(ntxent_b200_forward_rows(v.u_all, c->M, c->D, c->row_begin, c->rows, tau, c->op_dtype, c->flags, c->ws,
                                    c->ws_bytes, lse_mine, c->loss_part, st));
    // the loss is needed by nobody downstream: its all-reduce rides on the second stream
    NTX_CUDA(cudaEventRecord(c->ev_fork, st));
    NTX_CUDA(cudaStreamWaitEvent(c->side, c->ev_fork, 0));
    if (multi) NTX_NCCL(n.AllReduce(c->loss_part, c->loss_part + 1, 1, ncclFloat32, ncclSum, c->nccl_comm, c->side));
    NTX_CUDA(launch_kernel(loss_sum_kernel, dim3(1), dim3(32), 0, c->side,
                           (const float*)(multi ? c->loss_part + 1 : c->loss_part), 1, 1.0f / (float)c->M, loss_out));
    if (bwd_mode == NTXENT_BWD_SYMMETRIC) {
        if (multi) NTX_NCCL(n.AllGather(lse_mine, v.lse_all, (size_t)c->rows, ncclFloat32, c->nccl_comm, st));
        NTX_RC(backward_rows_dz_tickets(v.u_all, v.lse_all, c->M, c->D, c->row_begin, c->rows, tau, c->dtype,
                                        c->flags, c->ws, c->ws_bytes, z, c->inv_norm, g, dz_out, nullptr,
                                        nullptr, 0, nullptr, c->tickets + 2 + (c->rows + 127) / 128, st));
    } else {
        // contract form: column term of every global row first, its reduce-scatter on the second
        // stream while the row term is computed on the first
        NTX_RC(ntxent_b200_backward_rows_part(v.u_all, v.lse_all, c->M, c->D, c->row_begin, c->rows, tau, 1,
                                              c->op_dtype, c->flags, c->ws, c->ws_bytes, c->du_cols, st));
        NTX_CUDA(cudaEventRecord(c->ev_join, st));
        NTX_CUDA(cudaStreamWaitEvent(c->side, c->ev_join, 0));
        if (multi)
            NTX_NCCL(n.ReduceScatter(c->du_cols, c->du_col_local, (size_t)c->rows * c->D, ncclFloat32, ncclSum,
                                     c->nccl_comm, c->side));
        NTX_RC(ntxent_b200_backward_rows_part(v.u_all, v.lse_all, c->M, c->D, c->row_begin, c->rows, tau, 0,
                                              c->op_dtype, c->flags, c->ws, c->ws_bytes, c->du_rows, st));
        NTX_CUDA(cudaEventRecord(c->ev_join, c->side));
        NTX_CUDA(cudaStreamWaitEvent(st, c->ev_join, 0));
        const float* col = multi ? c->du_col_local : c->du_cols + (size_t)c->row_begin * c->D;
        const int64_t n4 = c->rows * c->D / 4;
        NTX_CUDA(launch_kernel(add_inplace_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0, st,
                               c->du_rows, col, n4));
        NTX_RC(ntxent_b200_finalize_dz(c->du_rows, z, c->inv_norm, g, c->rows, c->D, c->dtype, dz_out, st));
    }
    // join the second stream (loss) back into the caller's
    NTX_CUDA(cudaEventRecord(c->ev_join, c->side));
    NTX_CUDA(cudaStreamWaitEvent(st, c->ev_join, 0));
    return NTXENT_OK;
}

}  // namespace
}  // namespace ntxent_b200

extern "C" int ntxent_b200_dist_forward_backward(ntxent_comm_t* comm, const void* z_rows, const float* grad_out,
                                                 float tau, int exchange, int bwd_mode, float* loss_out,
                                                 void* dz_rows_out, ntxent_stream_t stream) {
    if (!comm || !z_rows || !loss_out || !dz_rows_out || !(tau > 0.f)) return NTXENT_ERR_INVALID_ARGUMENT;
    if (exchange != NTXENT_EXCHANGE_PUSH && exchange != NTXENT_EXCHANGE_NCCL) return NTXENT_ERR_INVALID_ARGUMENT;
    if (bwd_mode != NTXENT_BWD_SYMMETRIC && bwd_mode != NTXENT_BWD_CONTRACT) return NTXENT_ERR_INVALID_ARGUMENT;
    if (exchange == NTXENT_EXCHANGE_PUSH && bwd_mode != NTXENT_BWD_SYMMETRIC) return NTXENT_ERR_INVALID_ARGUMENT;
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != comm->dev) return NTXENT_ERR_INVALID_ARGUMENT;
    if (exchange == NTXENT_EXCHANGE_PUSH)
        return step_push(comm, z_rows, grad_out, tau, loss_out, dz_rows_out, stream);
    return step_nccl(comm, z_rows, grad_out, tau, bwd_mode, loss_out, dz_rows_out, stream);
}
