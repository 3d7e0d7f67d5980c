// capi.cu -- the extern "C" boundary declared in include/ntxent_b200.h: argument checks,
// workspace carving, kernel-family dispatch.  Nothing here allocates or synchronises.
#include "../../include/ntxent_b200.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

#include <nvtx3/nvToolsExt.h>  // header-only: ranges are no-ops unless a profiler injects itself

#include "common.cuh"
#include "tc_kernels.cuh"

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

int device_sm_count() {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) {
        (void)cudaGetLastError();
        return 148;
    }
    return sms;
}

bool pdl_enabled() {
    static const bool on = [] { const char* v = getenv("NTXENT_NO_PDL"); return !(v && v[0] == '1'); }();
    return on;
}

SplitPlan plan_splits(int64_t a_count, int64_t b_count, int row_tile, int col_tile, int slots,
                      int max_split) {
    SplitPlan best{};
    best.row_tiles = (int)((a_count + row_tile - 1) / row_tile);
    const int64_t col_tiles = std::max<int64_t>(1, (b_count + col_tile - 1) / col_tile);
    double best_cost = 1e300;
    const int lim = (int)std::min<int64_t>(max_split, col_tiles);
    for (int ns = 1; ns <= lim; ++ns) {
        const int64_t tiles_per = (col_tiles + ns - 1) / ns;
        const int64_t real_ns = (col_tiles + tiles_per - 1) / tiles_per;
        if (real_ns != ns) continue;
        const int64_t ctas = (int64_t)best.row_tiles * ns;
        const int64_t waves = (ctas + slots - 1) / slots;
        // per-CTA cost: its column tiles plus ~1.5 tiles of prologue/epilogue; a small
        // per-split charge for the partial that the epilogue has to re-read
        const double cost = (double)waves * ((double)tiles_per + 1.5) + 0.02 * ns;
        if (cost < best_cost) {
            best_cost = cost;
            best.nsplit = ns;
            best.cols_per_split = tiles_per * col_tile;
        }
    }
    return best;
}

namespace {

// Optional per-kernel timing for the benchmark (ntxent_b200_profile_begin/_end): CUDA events
// recorded on the launching stream around the two tile kernels.  Off by default; not
// thread-safe; the only place in the library that creates CUDA objects.
struct KernelTimer {
    bool on = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev[2];
    void start(int which, cudaStream_t st) {
        if (!on) return;
        cudaEvent_t a, b;
        if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
        ev[which].push_back({a, b});
        cudaEventRecord(a, st);
    }
    void stop(int which, cudaStream_t st) {
        if (!on || ev[which].empty()) return;
        cudaEventRecord(ev[which].back().second, st);
    }
};
KernelTimer g_timer;

constexpr int kMaxSplit = 32;
constexpr size_t kAlign = 256;
inline size_t align_up(size_t x) { return (x + kAlign - 1) / kAlign * kAlign; }

inline bool op_is_bf16(int dtype, unsigned flags) {
    return dtype == NTXENT_BF16 || (flags & NTXENT_FLAG_MIXED_PRECISION);
}
// tf32 tiles for fp32 operands only while log2(e)/(tau sqrt(D)) <= 1.7 (tau >= 0.075 at D = 128,
// >= 0.106 at D = 64): their error in dz grows with 1/tau and shrinks with sqrt(D) (measured against
// the fp64 oracle: 6e-5 at D 128 / tau 0.07, 1.2e-4..2.4e-4 at D 64 / tau 0.07, bound 1e-4); colder
// than that the exact fp32-FMA kernels run.  tau < 0 means "not known yet" (workspace sizing).
constexpr float kTf32MaxScaledK = 1.7f;
inline bool want_tc(int64_t M, int64_t D, bool op_bf16, unsigned flags, float tau) {
    if (flags & NTXENT_FLAG_FORCE_SIMT) return false;
    // the cross-view form walks whole column tiles of the other view: N = M/2 must be tile-aligned
    if ((flags & NTXENT_FLAG_CROSS_VIEW) && (M / 2) % 128 != 0) return false;
    if (!op_bf16 && tau >= 0.f && kLog2e / (tau * sqrtf((float)D)) > kTf32MaxScaledK) return false;
    return tc_supported(M, D, op_bf16);
}

// Internal flag (never accepted from callers): keep the whole-batch forward on the full-matrix
// tile kernel -- for ntxent_b200_forward_static, whose built-in scratch has no room for the
// upper-triangle kernel's column partials.
constexpr unsigned kFlagNoSymForward = 1u << 16;
constexpr unsigned kPublicFlags = NTXENT_FLAG_MIXED_PRECISION | NTXENT_FLAG_FORCE_SIMT | NTXENT_FLAG_ONLINE_MAX |
                                  NTXENT_FLAG_CROSS_VIEW;

struct Plans {
    SplitPlan fwd, bwd_rows, bwd_cols;
    bool tc;
    unsigned flags;
    size_t ut_bytes;
    // whole-batch forward over the upper triangle (tc_fwd_sym.cu); `sym_room`: the workspace
    // layout reserves its buffers (decided without tau), `sym`: this call uses it
    bool sym_room, sym;
    SymPlan symp;
    bool cross;  // NTXENT_FLAG_CROSS_VIEW: denominators over the other view only (whole-batch calls)
};

// All split decisions in one place so that sizing and launching can never disagree.
Plans make_plans(int64_t M, int64_t D, int64_t row_count, bool op_bf16, unsigned flags, float tau) {
    Plans p{};
    p.flags = flags;
    p.ut_bytes = 0;
    p.tc = want_tc(M, D, op_bf16, flags, tau);
    p.cross = (flags & NTXENT_FLAG_CROSS_VIEW) != 0;
    const int sms = device_sm_count();
    if (p.tc) {
        // (cross-view form: every row tile meets only the M/2 columns of the other view)
        p.fwd = tc_plan_fwd(p.cross ? M / 2 : M, D, row_count, op_bf16, sms);
        p.bwd_rows = tc_plan_bwd(M, D, row_count, p.cross ? M / 2 : M, sms);
        p.bwd_cols = tc_plan_bwd(M, D, M, row_count, sms);
        p.ut_bytes = tc_bwd_scratch_bytes(M, D, op_bf16);
        const bool env_off = [] { const char* v = getenv("NTXENT_NO_SYM_FORWARD"); return v && v[0] == '1'; }();
        if (row_count == M && !(flags & kFlagNoSymForward) && !env_off && !p.cross) {
            p.sym_room = tc_sym_supported(M, D, op_bf16, 1.0f, flags);
            p.sym = tau >= 0.f && tc_sym_supported(M, D, op_bf16, tau, flags);
            if (p.sym_room) p.symp = tc_sym_plan(M, D, sms);
        }
    } else {
        p.fwd = plan_splits(row_count, M, 64, 64, sms * 4, kMaxSplit);
        p.bwd_rows = plan_splits(row_count, M, 64, 64, sms * 2, kMaxSplit);
        p.bwd_cols = plan_splits(M, row_count, 64, 64, sms * 2, kMaxSplit);
    }
    return p;
}

// Tickets: [0] the loss (fused forward finalize), [1, 1+T) the in-kernel forward finalize, one per
// 128-row tile, [1+T, 1+2T) the in-kernel backward epilogue (bwd_tickets).
inline size_t ticket_words(int64_t row_count) { return 1 + 2 * (size_t)((row_count + 127) / 128); }
inline unsigned* bwd_tickets(unsigned* counter, int64_t row_count) {
    return counter + 1 + (size_t)((row_count + 127) / 128);
}
// The backward epilogue CAN run inside the tile kernel (NTXENT_BWD_TAIL=1) when that kernel has
// the code for it (bf16 operands, D <= 128), nothing but dz is asked for, and a row tile is shared
// by at most two CTAs: then the epilogues are spread over the kernel's run.  Measured at C4 on one
// GPU, interleaved on one box: step 0.6404 against 0.6476 ms (-1.1 %), tile kernel +10 us, side
// kernels -17 us.  Off by default: the gain is inside the box-to-box spread (+-2 %), it exists only
// for that shape class (row blocks that many CTAs share finish all at once at the end of the kernel,
// where the separate, fully parallel epilogue kernel is the faster one: 4096 rows 29.5 against
// 32.8 us), and it keeps a second copy of the epilogue arithmetic that must stay bit-identical
// (tests/test_parity_gpu.py pins it).  Read per call so that a test can switch it.
inline bool bwd_tail_enabled(const Plans& p, int64_t D, bool op_bf16) {
    const char* v = getenv("NTXENT_BWD_TAIL");
    return v && v[0] == '1' && p.tc && tc_bwd_tail_supported(D, op_bf16) && p.bwd_rows.sk_ctas > 0 &&
           p.bwd_rows.nsplit <= 2;
}

struct RowsWs {
    float* pos_t;  // positive logits (base-2 units): captured by the upper-triangle forward, or dot products
    float* block_partials;
    unsigned* counter;  // ticket of the fused forward finalize (whole-batch calls: zeroed by normalise)
    float* part_m;
    float* part_s;
    float* acc_rows;
    float* acc_cols;
    float* ut;  // tf32 tile kernels: transposed copy of U
    float* sym_part_s;  // upper-triangle forward: row-sum partials [slots][M] ...
    float* sym_cpart;   // ... and column partials
    size_t bytes;
};

RowsWs carve_rows(char* base, int64_t M, int64_t D, int64_t row_count, const Plans& p,
                  bool with_cols) {
    RowsWs w{};
    size_t off = 0;
    auto take = [&](size_t n) {
        char* ptr = base ? base + off : nullptr;
        off += align_up(n);
        return ptr;
    };
    w.pos_t = (float*)take(sizeof(float) * (size_t)row_count);
    w.block_partials = (float*)take(sizeof(float) * (size_t)((row_count + 7) / 8));
    w.counter = (unsigned*)take(sizeof(unsigned) * ticket_words(row_count));
    w.part_m = (float*)take(sizeof(float) * (size_t)p.fwd.nsplit * row_count);
    w.part_s = (float*)take(sizeof(float) * (size_t)p.fwd.nsplit * row_count);
    w.acc_rows = (float*)take(sizeof(float) * (size_t)p.bwd_rows.nsplit * row_count * D);
    if (with_cols) w.acc_cols = (float*)take(sizeof(float) * (size_t)p.bwd_cols.nsplit * M * D);
    if (p.ut_bytes) w.ut = (float*)take(p.ut_bytes);
    if (p.sym_room) {  // (last, so that nothing in front of them moves)
        w.sym_part_s = (float*)take(sizeof(float) * (size_t)p.symp.slots * M);
        w.sym_cpart = (float*)take(sizeof(float) * p.symp.cpart_floats);
    }
    w.bytes = off;
    return w;
}

struct DeviceWs {
    void* U;
    float* inv_norm;
    RowsWs rows;
    size_t bytes;
};

DeviceWs carve_device(char* base, int64_t M, int64_t D, bool op_bf16, const Plans& p) {
    DeviceWs w{};
    size_t off = 0;
    w.U = base;
    off += align_up((size_t)M * D * (op_bf16 ? 2 : 4));
    w.inv_norm = (float*)(base ? base + off : nullptr);
    off += align_up(sizeof(float) * (size_t)M);
    w.rows = carve_rows(base ? base + off : nullptr, M, D, M, p, false);
    w.bytes = off + w.rows.bytes;
    return w;
}

// The kernels exist for sm_100 only: say so instead of letting the first launch fail with
// "no kernel image".  One attribute query per device, remembered.
int check_device() {
    static std::atomic<unsigned long long> ok_mask{0};
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return NTXENT_ERR_UNSUPPORTED_DEVICE; }
    if (dev < 64 && ((ok_mask.load(std::memory_order_relaxed) >> dev) & 1ull)) return NTXENT_OK;
    if (cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
        (void)cudaGetLastError();
        return NTXENT_ERR_UNSUPPORTED_DEVICE;
    }
    if (major != 10) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    if (dev < 64) ok_mask.fetch_or(1ull << dev, std::memory_order_relaxed);
    return NTXENT_OK;
}

int check_shape(int64_t M, int64_t D, float tau, int dtype) {
    // (flags: unknown bits are ignored, as before; the internal ones are masked off at the boundary)
    if (dtype != NTXENT_F32 && dtype != NTXENT_BF16) return NTXENT_ERR_INVALID_ARGUMENT;
    if (M < 2 || (M & 1) || !(tau > 0.f) || !(tau < INFINITY)) return NTXENT_ERR_INVALID_ARGUMENT;
    if (D < 8 || (D % 8) != 0 || D > 1024) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    if (M > (int64_t)1 << 24) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    return NTXENT_OK;
}

// NVTX ranges around the compute entry points (the reference's ENABLE_PROFILING build option,
// CMakeLists.txt:10,82-84): on with NTXENT_NVTX=1, for timelines in a profiler; off by default so
// that a step costs no host time for them.
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char* name) {
        static const bool enabled = [] { const char* v = getenv("NTXENT_NVTX"); return v && v[0] == '1'; }();
        on = enabled;
        if (on) nvtxRangePushA(name);
    }
    ~NvtxRange() { if (on) nvtxRangePop(); }
};
#define NTX_RANGE(name) NvtxRange nvtx_range__(name)

// NTXENT_DEBUG=1 in the environment names the failing launch on stderr.
inline bool debug_on() {
    static const bool on = [] { const char* v = getenv("NTXENT_DEBUG"); return v && v[0] == '1'; }();
    return on;
}
// TMA and the vector loads want 16-byte aligned rows; workspaces are carved in 256-byte units.
inline bool aligned(const void* p, uintptr_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }

#define NTX_TRY(expr)                                                                         \
    do {                                                                                      \
        cudaError_t e__ = (expr);                                                             \
        if (e__ != cudaSuccess) {                                                             \
            if (debug_on())                                                                   \
                fprintf(stderr, "[ntxent_b200] %s:%d: %s -> %s\n", __FILE__, __LINE__, #expr, \
                        cudaGetErrorString(e__));                                             \
            return (int)e__;                                                                  \
        }                                                                                     \
    } while (0)

int run_forward_rows_push(const void* u_all, bool op_bf16, int64_t M, int64_t D, int64_t row_begin,
                          int64_t row_count, float tau, const Plans& p, const RowsWs& w,
                          const PeerList& lse_outs, int64_t lse_offset, const PeerList& loss_outs,
                          int loss_index, float loss_scale, const PeerWait& u_ready,
                          const PeerSignal& lse_signal, cudaStream_t st, const int* pos_index = nullptr,
                          unsigned* counter = nullptr) {
    g_timer.start(0, st);
    if (p.tc) {
        // With tickets the tile kernel CAN finalize in its own tail (no further launch).  Measured
        // slower than the finalize kernel it replaces -- the tail's chain of latencies (partials from
        // L2, fence, ticket, peer stores, system fence, ticket, sum, flag) runs on a handful of CTAs at
        // the very end of the kernel: +20 us on the tile kernel for -11 us of launches at a 4096-row
        // shard (profiles/r2_fusion_ab.txt) -- so it is off unless NTXENT_FWD_TAIL=1.
        static const bool use_tail = [] { const char* v = getenv("NTXENT_FWD_TAIL"); return v && v[0] == '1'; }();
        unsigned* const tail_tickets = use_tail ? counter : nullptr;
        FwdTail tail{};
        tail.tickets = tail_tickets;
        tail.pos_index = pos_index;
        tail.lse_outs = lse_outs; tail.lse_offset = lse_offset;
        tail.block_partials = w.block_partials;
        tail.loss_outs = loss_outs; tail.loss_index = loss_index; tail.loss_scale = loss_scale;
        tail.signal = lse_signal;
        NTX_TRY(tc_launch_fwd(u_all, op_bf16, M, D, row_begin, row_count, tau, p.flags, p.fwd, w.part_m,
                              w.part_s, u_ready, st, p.cross, tail_tickets ? &tail : nullptr));
        g_timer.stop(0, st);
        if (tail_tickets) return NTXENT_OK;
    } else {
        NTX_TRY(launch_simt_fwd(u_all, op_bf16, M, D, row_begin, row_count, tau, p.fwd, w.part_m,
                                w.part_s, u_ready, st, p.cross ? M / 2 : 0));
        g_timer.stop(0, st);
    }
    NTX_TRY(launch_fwd_finalize_push(u_all, op_bf16, M, D, row_begin, row_count, tau, p.fwd.nsplit, w.part_m,
                                     w.part_s, w.pos_t, pos_index, lse_outs, lse_offset, w.block_partials,
                                     loss_outs, loss_index, loss_scale, lse_signal, st, counter));
    return NTXENT_OK;
}

int run_forward_rows(const void* u_all, bool op_bf16, int64_t M, int64_t D, int64_t row_begin,
                     int64_t row_count, float tau, const Plans& p, const RowsWs& w, float* lse_rows,
                     float* loss_out, float loss_scale, cudaStream_t st, const int* pos_index = nullptr,
                     unsigned* counter = nullptr) {
    return run_forward_rows_push(u_all, op_bf16, M, D, row_begin, row_count, tau, p, w,
                                 single_peer(lse_rows), 0, single_peer(loss_out), 0, loss_scale, no_wait(),
                                 no_signal(), st, pos_index, counter);
}

// Forward of the whole batch on one device: over the upper triangle where that applies
// (tc_fwd_sym.cu), else the row-block path with the full range.
int run_forward_device(const void* U, bool op_bf16, int64_t M, int64_t D, float tau, const Plans& p,
                       const RowsWs& w, float* lse_out, float* loss_out, cudaStream_t st,
                       const int* pos_index = nullptr) {
    if (!p.sym)
        return run_forward_rows(U, op_bf16, M, D, 0, M, tau, p, w, lse_out, loss_out, 1.0f / (float)M, st,
                                pos_index, w.counter);
    g_timer.start(0, st);
    NTX_TRY(tc_launch_fwd_sym(U, M, D, tau, p.symp, w.sym_part_s, w.sym_cpart, w.pos_t, st));
    g_timer.stop(0, st);
    if (pos_index) NTX_TRY(launch_pos_dot(U, op_bf16, M, D, 0, M, pos_index, kLog2e / tau, w.pos_t, st));
    NTX_TRY(tc_launch_fwd_sym_finalize(M, p.symp, w.sym_part_s, w.sym_cpart, w.pos_t, lse_out, w.block_partials, st,
                                       w.counter, loss_out, 1.0f / (float)M));
    return NTXENT_OK;
}

bool to_peer_signal(const ntxent_peer_signal* in, PeerSignal* out);
bool to_peer_wait(const ntxent_peer_wait* in, PeerWait* out);

bool to_peer_list(const ntxent_peer_list* in, PeerList* out) {
    if (!in || in->count < 1 || in->count > NTXENT_MAX_PEERS || in->count > kMaxPeers) return false;
    out->n = in->count;
    for (int i = 0; i < in->count; ++i) {
        if (!in->ptr[i] || !aligned(in->ptr[i], 16)) return false;
        out->p[i] = in->ptr[i];
    }
    return true;
}

// NULL selects "off"; anything else must be complete.
bool to_peer_signal(const ntxent_peer_signal* in, PeerSignal* out) {
    *out = no_signal();
    if (!in) return true;
    if (!in->epoch || in->rank < 0 || in->rank >= kMaxPeers ||
        !to_peer_list(&in->flags, &out->flags))
        return false;
    out->epoch = in->epoch;
    out->rank = in->rank;
    out->ticket = in->ticket;
    return true;
}
// How long a kernel waits for a peer's flag before it gives up (ntxent_b200_set_peer_wait_timeout_ms).
std::atomic<unsigned long long> g_peer_timeout_ns{600ull * 1000ull * 1000ull * 1000ull};

bool to_peer_wait(const ntxent_peer_wait* in, PeerWait* out) {
    *out = no_wait();
    if (!in) return true;
    if (!in->flags || !in->epoch || in->ranks < 1 || in->ranks > kMaxPeers || in->rows_per_rank < 1)
        return false;
    out->flags = in->flags;
    out->epoch = in->epoch;
    out->ranks = in->ranks;
    out->rows_per_rank = in->rows_per_rank;
    out->timeout_ns = g_peer_timeout_ns.load(std::memory_order_relaxed);
    out->status = in->status;
    return true;
}

int run_bwd_tiles(const void* u_all, bool op_bf16, const float* lse_all, int64_t M, int64_t D,
                  int64_t a_begin, int64_t a_count, int64_t b_begin, int64_t b_count, float tau,
                  int mode, const Plans& p, const SplitPlan& plan, float* acc, float* ut,
                  cudaStream_t st, const PeerWait& lse_ready = PeerWait{}, const BwdTail* tail = nullptr) {
    g_timer.start(1, st);
    if (p.tc)
        NTX_TRY(tc_launch_bwd(u_all, op_bf16, lse_all, M, D, a_begin, a_count, b_begin, p.cross ? M / 2 : b_count,
                              tau, mode, p.flags, plan, acc, ut, lse_ready, st, p.cross, tail));
    else
        NTX_TRY(launch_simt_bwd(u_all, op_bf16, lse_all, M, D, a_begin, a_count, b_begin, b_count,
                                tau, mode, plan, acc, lse_ready, st, p.cross ? M / 2 : 0));
    g_timer.stop(1, st);
    return NTXENT_OK;
}

// Row block I = [row_begin, row_begin + row_count) against all columns: the symmetric dU of the
// block (sym) or the row term G_{I,:} U of the contract form -> du_rows_out [row_count, D].
int run_bwd_row_term(const void* u_all, bool opb, const float* lse_all, int64_t M, int64_t D, int64_t row_begin,
                     int64_t row_count, float tau, bool sym, const Plans& p, const RowsWs& w, float* du_rows_out,
                     cudaStream_t stream) {
    int rc = run_bwd_tiles(u_all, opb, lse_all, M, D, row_begin, row_count, 0, M, tau, sym ? kSym : kRow, p,
                           p.bwd_rows, w.acc_rows, w.ut, stream);
    if (rc != NTXENT_OK) return rc;
    BwdEpilogueArgs e{};
    e.acc = w.acc_rows; e.nsplit = p.bwd_rows.nsplit;
    e.sk_ctas = p.bwd_rows.sk_ctas; e.sk_col_tiles = p.bwd_rows.sk_col_tiles;
    e.u_all = u_all; e.op_bf16 = opb;
    e.M = M; e.D = D; e.a_begin = row_begin; e.a_count = row_count;
    e.pos_coeff = sym ? 2.f : 1.f; e.pos_lo = 0; e.pos_hi = M;
    e.scale = 1.0f / (tau * (float)M);
    e.du_out = du_rows_out;
    NTX_TRY(launch_bwd_epilogue(e, stream));
    return NTXENT_OK;
}
// All rows against the columns of block I: the column term G_{I,:}^T U_I -> du_cols_out [M, D].
int run_bwd_col_term(const void* u_all, bool opb, const float* lse_all, int64_t M, int64_t D, int64_t row_begin,
                     int64_t row_count, float tau, const Plans& p, const RowsWs& w, float* du_cols_out,
                     cudaStream_t stream) {
    int rc = run_bwd_tiles(u_all, opb, lse_all, M, D, 0, M, row_begin, row_count, tau, kCol, p, p.bwd_cols,
                           w.acc_cols, w.ut, stream);
    if (rc != NTXENT_OK) return rc;
    BwdEpilogueArgs c{};
    c.acc = w.acc_cols; c.nsplit = p.bwd_cols.nsplit;
    c.sk_ctas = p.bwd_cols.sk_ctas; c.sk_col_tiles = p.bwd_cols.sk_col_tiles;
    c.u_all = u_all; c.op_bf16 = opb;
    c.M = M; c.D = D; c.a_begin = 0; c.a_count = M;
    c.pos_coeff = 1.f; c.pos_lo = row_begin; c.pos_hi = row_begin + row_count;
    c.scale = 1.0f / (tau * (float)M);
    c.du_out = du_cols_out;
    NTX_TRY(launch_bwd_epilogue(c, stream));
    return NTXENT_OK;
}

int run_backward_device(const void* z, bool z_bf16, const DeviceWs& w, bool op_bf16, const float* lse,
                        const float* grad_out, int64_t M, int64_t D, float tau, const Plans& p,
                        void* dz_out, float* pos_prob_out, cudaStream_t st,
                        float* tau_grad_rows_out = nullptr, const int* pos_index = nullptr,
                        const int* inv_ptr = nullptr, const int* inv_idx = nullptr) {
    if (bwd_tail_enabled(p, D, op_bf16) && !pos_prob_out && !tau_grad_rows_out && !pos_index) {
        BwdTail t{};
        t.tickets = bwd_tickets(w.rows.counter, M);  // (zeroed by this call's normalise launch)
        t.z_rows = z; t.z_bf16 = z_bf16; t.inv_norm = w.inv_norm; t.grad_out = grad_out; t.dz_out = dz_out;
        t.pos_coeff = 2.f; t.scale = 1.0f / (tau * (float)M);
        return run_bwd_tiles(w.U, op_bf16, lse, M, D, 0, M, 0, M, tau, kSym, p, p.bwd_rows, w.rows.acc_rows,
                             w.rows.ut, st, PeerWait{}, &t);
    }
    int rc = run_bwd_tiles(w.U, op_bf16, lse, M, D, 0, M, 0, M, tau, kSym, p, p.bwd_rows,
                           w.rows.acc_rows, w.rows.ut, st);
    if (rc != NTXENT_OK) return rc;
    BwdEpilogueArgs e{};
    e.acc = w.rows.acc_rows; e.nsplit = p.bwd_rows.nsplit;
    e.sk_ctas = p.bwd_rows.sk_ctas; e.sk_col_tiles = p.bwd_rows.sk_col_tiles;
    e.u_all = w.U; e.op_bf16 = op_bf16;
    e.M = M; e.D = D; e.a_begin = 0; e.a_count = M;
    e.pos_coeff = 2.f; e.pos_lo = 0; e.pos_hi = M;
    e.scale = 1.0f / (tau * (float)M);
    e.du_out = nullptr; e.z_rows = z; e.z_bf16 = z_bf16; e.inv_norm = w.inv_norm;
    e.grad_out = grad_out; e.dz_out = dz_out;
    e.lse_all = lse; e.inv_tau = 1.0f / tau; e.pos_prob_out = pos_prob_out;
    e.tau_grad_rows_out = tau_grad_rows_out;
    e.pos_index = pos_index; e.inv_ptr = inv_ptr; e.inv_idx = inv_idx;
    NTX_TRY(launch_bwd_epilogue(e, st));
    return NTXENT_OK;
}

}  // namespace
}  // namespace ntxent_b200

using namespace ntxent_b200;

extern "C" {

int ntxent_b200_device_supported(void) {
    int dev = 0, major = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) {
        (void)cudaGetLastError();
        return 0;
    }
    return major == 10 ? 1 : 0;
}

const char* ntxent_b200_version(void) { return "ntxent_b200 0.1 (sm_100a)"; }

const char* ntxent_b200_strerror(int code) {
    switch (code) {
        case NTXENT_OK: return "ok";
        case NTXENT_ERR_INVALID_ARGUMENT:
            return "invalid argument (null or misaligned pointer, odd or too small M, tau <= 0, unknown "
                   "dtype/mode; tensors need 16-byte, workspaces 256-byte alignment)";
        case NTXENT_ERR_UNSUPPORTED_DEVICE: return "device is not sm_100 (B200)";
        case NTXENT_ERR_WORKSPACE_TOO_SMALL:
            return "workspace smaller than ntxent_b200_*workspace_bytes()";
        case NTXENT_ERR_UNSUPPORTED_SHAPE:
            return "unsupported shape (need 8 <= D <= 1024, D % 8 == 0, M <= 2^24)";
        case NTXENT_ERR_STATIC_SCRATCH_EXCEEDED:
            return "problem needs more scratch than the built-in NTXENT_STATIC_SCRATCH_BYTES array "
                   "of ntxent_b200_forward_static; use ntxent_b200_forward with a caller workspace";
        case NTXENT_ERR_COLLECTIVE:
            return "collective library: libnccl.so.2 could not be loaded, the communicator was made without "
                   "it, or a NCCL call failed (NTXENT_DEBUG=1 names the call)";
        case NTXENT_ERR_NO_PEER_ACCESS:
            return "the devices of the communicator cannot map each other's memory (no peer access / CUDA IPC)";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown ntxent_b200 error";
}

int ntxent_b200_profile_begin(void) {
    g_timer.on = true;
    return NTXENT_OK;
}

int ntxent_b200_profile_end(float* fwd_ms, int* fwd_launches, float* bwd_ms, int* bwd_launches) {
    g_timer.on = false;
    float* ms[2] = {fwd_ms, bwd_ms};
    int* cnt[2] = {fwd_launches, bwd_launches};
    int rc = NTXENT_OK;
    for (int k = 0; k < 2; ++k) {
        float total = 0.f;
        int n = 0;
        for (auto& pr : g_timer.ev[k]) {
            float t = 0.f;
            cudaError_t e = cudaEventSynchronize(pr.second);
            if (e == cudaSuccess) e = cudaEventElapsedTime(&t, pr.first, pr.second);
            if (e == cudaSuccess) { total += t; ++n; } else rc = (int)e;
            cudaEventDestroy(pr.first);
            cudaEventDestroy(pr.second);
        }
        g_timer.ev[k].clear();
        if (ms[k]) *ms[k] = total;
        if (cnt[k]) *cnt[k] = n;
    }
    return rc;
}

int ntxent_b200_debug_set_trace(void* dev_buffer) {
    return tc_set_trace_buffer((unsigned long long*)dev_buffer) ? NTXENT_OK : NTXENT_ERR_INVALID_ARGUMENT;
}

int ntxent_b200_plan_info(int64_t M, int64_t D, int64_t row_count, int dtype, unsigned flags,
                          int* uses_tcgen05, int* fwd_splits, int* bwd_splits) {
    if (check_shape(M, D, 1.0f, dtype) != NTXENT_OK || row_count < 1 || row_count > M)
        return NTXENT_ERR_INVALID_ARGUMENT;
    const Plans p = make_plans(M, D, row_count, op_is_bf16(dtype, flags), flags, -1.f);
    if (uses_tcgen05) *uses_tcgen05 = p.tc ? 1 : 0;
    if (fwd_splits) *fwd_splits = p.fwd.nsplit;
    if (bwd_splits) *bwd_splits = p.bwd_rows.nsplit;
    return NTXENT_OK;
}

int ntxent_b200_operand_size(int dtype, unsigned flags) { return op_is_bf16(dtype, flags) ? 2 : 4; }

size_t ntxent_b200_workspace_bytes(int64_t M, int64_t D, int dtype, unsigned flags) {
    if (check_shape(M, D, 1.0f, dtype) != NTXENT_OK) return 0;
    const bool opb = op_is_bf16(dtype, flags);
    // fp32 operands may run on either kernel family depending on tau: size for the larger one
    const size_t a = carve_device(nullptr, M, D, opb, make_plans(M, D, M, opb, flags, -1.f)).bytes;
    const size_t b = carve_device(nullptr, M, D, opb,
                                  make_plans(M, D, M, opb, flags | NTXENT_FLAG_FORCE_SIMT, -1.f)).bytes;
    return std::max(a, b);
}

size_t ntxent_b200_rows_workspace_bytes(int64_t M, int64_t D, int64_t row_count, int dtype,
                                        unsigned flags) {
    if (check_shape(M, D, 1.0f, dtype) != NTXENT_OK || row_count < 1 || row_count > M) return 0;
    const bool opb = op_is_bf16(dtype, flags);
    const size_t a = carve_rows(nullptr, M, D, row_count, make_plans(M, D, row_count, opb, flags, -1.f), true).bytes;
    const size_t b = carve_rows(nullptr, M, D, row_count,
                                make_plans(M, D, row_count, opb, flags | NTXENT_FLAG_FORCE_SIMT, -1.f), true)
                         .bytes;
    return std::max(a, b);
}

int ntxent_b200_normalize(const void* z, int64_t rows, int64_t D, int dtype, unsigned flags,
                          void* u_out, float* inv_norm_out, ntxent_stream_t stream) {
    if (!z || !u_out || rows < 0 || !aligned(z, 16) || !aligned(u_out, 16)) return NTXENT_ERR_INVALID_ARGUMENT;
    if (dtype != NTXENT_F32 && dtype != NTXENT_BF16) return NTXENT_ERR_INVALID_ARGUMENT;
    if (D < 8 || (D % 8) != 0 || D > 1024) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, rows, D, u_out, op_is_bf16(dtype, flags),
                             inv_norm_out, stream));
    return NTXENT_OK;
}

int ntxent_b200_normalize_push(const void* z, int64_t rows, int64_t D, int dtype, unsigned flags,
                               const ntxent_peer_list* u_all_peers, int64_t row_begin,
                               float* inv_norm_out, const ntxent_peer_signal* u_signal,
                               ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_normalize_push");
    PeerList outs{};
    PeerSignal sig{};
    if (!z || rows < 0 || row_begin < 0 || !aligned(z, 16) || !to_peer_list(u_all_peers, &outs) ||
        !to_peer_signal(u_signal, &sig) || (u_signal && rows < 1))
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (dtype != NTXENT_F32 && dtype != NTXENT_BF16) return NTXENT_ERR_INVALID_ARGUMENT;
    if (D < 8 || (D % 8) != 0 || D > 1024) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize_push(z, dtype == NTXENT_BF16, rows, D, outs, row_begin,
                                  op_is_bf16(dtype, flags), inv_norm_out, sig, stream));
    return NTXENT_OK;
}

int ntxent_b200_forward_rows_push(const void* u_all, int64_t M, int64_t D, int64_t row_begin,
                                  int64_t row_count, float tau, int dtype, unsigned flags,
                                  void* workspace, size_t workspace_bytes,
                                  const ntxent_peer_list* lse_all_peers,
                                  const ntxent_peer_list* loss_parts_peers, int loss_index,
                                  const ntxent_peer_wait* u_ready, const ntxent_peer_signal* lse_signal,
                                  ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_forward_rows_push");
    PeerList lse_outs{}, loss_outs{};
    PeerWait wait{};
    PeerSignal sig{};
    if (!u_all || !workspace || !to_peer_list(lse_all_peers, &lse_outs) ||
        !to_peer_list(loss_parts_peers, &loss_outs) || loss_index < 0 || !to_peer_wait(u_ready, &wait) ||
        !to_peer_signal(lse_signal, &sig) || (lse_signal && sig.flags.n != loss_outs.n))
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(u_all, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    if (row_begin < 0 || row_count < 1 || row_begin + row_count > M)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (flags & NTXENT_FLAG_CROSS_VIEW) return NTXENT_ERR_INVALID_ARGUMENT;  // whole-batch calls only
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, row_count, opb, flags, tau);
    const RowsWs w = carve_rows((char*)workspace, M, D, row_count, p, true);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    return run_forward_rows_push(u_all, opb, M, D, row_begin, row_count, tau, p, w, lse_outs, row_begin,
                                 loss_outs, loss_index, 1.0f, wait, sig, stream, nullptr, sig.ticket);
}

int ntxent_b200_set_peer_wait_timeout_ms(uint64_t milliseconds) {
    g_peer_timeout_ns.store((unsigned long long)milliseconds * 1000000ull, std::memory_order_relaxed);
    return NTXENT_OK;
}

int ntxent_b200_peer_wait(const ntxent_peer_wait* ready, ntxent_stream_t stream) {
    PeerWait wait{};
    if (!ready || !to_peer_wait(ready, &wait)) return NTXENT_ERR_INVALID_ARGUMENT;
    NTX_TRY(launch_peer_wait(wait, stream));
    return NTXENT_OK;
}

int ntxent_b200_backward_rows_dz(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                                 int64_t row_begin, int64_t row_count, float tau, int dtype,
                                 unsigned flags, void* workspace, size_t workspace_bytes,
                                 const void* z_rows, const float* inv_norm_rows, const float* grad_out,
                                 void* dz_out, const ntxent_peer_wait* lse_ready,
                                 const float* loss_parts, int loss_parts_count, float* loss_out,
                                 ntxent_stream_t stream) {
    return ntxent_b200::backward_rows_dz_tickets(u_all, lse_all, M, D, row_begin, row_count, tau, dtype, flags,
                                                 workspace, workspace_bytes, z_rows, inv_norm_rows, grad_out, dz_out,
                                                 lse_ready, loss_parts, loss_parts_count, loss_out, nullptr, stream);
}

// The same with a ticket array for the in-kernel epilogue (one zeroed word per 128-row tile that
// stays with the caller between calls: the communicator has one; the public entry point, whose
// workspace may hold anything, passes none and gets the epilogue kernel).
extern "C++" {
int ntxent_b200::backward_rows_dz_tickets(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                                          int64_t row_begin, int64_t row_count, float tau, int dtype,
                                          unsigned flags, void* workspace, size_t workspace_bytes,
                                          const void* z_rows, const float* inv_norm_rows, const float* grad_out,
                                          void* dz_out, const ntxent_peer_wait* lse_ready,
                                          const float* loss_parts, int loss_parts_count, float* loss_out,
                                          unsigned* tickets, cudaStream_t stream) {
    NTX_RANGE("ntxent_b200_backward_rows_dz");
    PeerWait wait{};
    if (!u_all || !lse_all || !workspace || !z_rows || !inv_norm_rows || !dz_out ||
        !to_peer_wait(lse_ready, &wait))
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(u_all, 16) || !aligned(workspace, 256) || !aligned(z_rows, 16) || !aligned(dz_out, 16))
        return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    if (row_begin < 0 || row_count < 1 || row_begin + row_count > M)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (flags & NTXENT_FLAG_CROSS_VIEW) return NTXENT_ERR_INVALID_ARGUMENT;  // whole-batch calls only
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, row_count, opb, flags, tau);
    const RowsWs w = carve_rows((char*)workspace, M, D, row_count, p, true);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    if (loss_out && (!loss_parts || loss_parts_count < 1 || loss_parts_count > kMaxPeers))
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (tickets && bwd_tail_enabled(p, D, opb)) {
        BwdTail t{};
        t.tickets = tickets;
        t.z_rows = z_rows; t.z_bf16 = (dtype == NTXENT_BF16); t.inv_norm = inv_norm_rows; t.grad_out = grad_out;
        t.dz_out = dz_out; t.pos_coeff = 2.f; t.scale = 1.0f / (tau * (float)M);
        if (loss_out) {
            t.loss_parts = loss_parts; t.loss_n = loss_parts_count; t.loss_scale = 1.0f / (float)M; t.loss_out = loss_out;
        }
        return run_bwd_tiles(u_all, opb, lse_all, M, D, row_begin, row_count, 0, M, tau, kSym, p, p.bwd_rows,
                             w.acc_rows, w.ut, stream, wait, &t);
    }
    rc = run_bwd_tiles(u_all, opb, lse_all, M, D, row_begin, row_count, 0, M, tau, kSym, p, p.bwd_rows,
                       w.acc_rows, w.ut, stream, wait);
    if (rc != NTXENT_OK) return rc;
    BwdEpilogueArgs e{};
    e.acc = w.acc_rows; e.nsplit = p.bwd_rows.nsplit;
    e.sk_ctas = p.bwd_rows.sk_ctas; e.sk_col_tiles = p.bwd_rows.sk_col_tiles;
    e.u_all = u_all; e.op_bf16 = opb;
    e.M = M; e.D = D; e.a_begin = row_begin; e.a_count = row_count;
    e.pos_coeff = 2.f; e.pos_lo = 0; e.pos_hi = M;
    e.scale = 1.0f / (tau * (float)M);
    e.du_out = nullptr; e.z_rows = z_rows; e.z_bf16 = (dtype == NTXENT_BF16); e.inv_norm = inv_norm_rows;
    e.grad_out = grad_out; e.dz_out = dz_out;
    if (loss_out) {
        e.loss_parts = loss_parts; e.loss_n = loss_parts_count; e.loss_scale = 1.0f / (float)M; e.loss_out = loss_out;
    }
    NTX_TRY(launch_bwd_epilogue(e, stream));
    return NTXENT_OK;
}
}  // extern "C++"

int ntxent_b200_forward(const void* z, int64_t M, int64_t D, float tau, int dtype, unsigned flags,
                        void* workspace, size_t workspace_bytes, float* loss_out, float* lse_out,
                        ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_forward");
    if (!z || !workspace || !loss_out || !lse_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(z, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, M, opb, flags, tau);
    const DeviceWs w = carve_device((char*)workspace, M, D, opb, p);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, M, D, w.U, opb, w.inv_norm, stream, w.rows.counter,
                             (int)ticket_words(M)));
    return run_forward_device(w.U, opb, M, D, tau, p, w.rows, lse_out, loss_out, stream);
}

int ntxent_b200_backward(const void* z, const float* lse, const float* grad_out, int64_t M,
                         int64_t D, float tau, int dtype, unsigned flags, void* workspace,
                         size_t workspace_bytes, void* dz_out, float* pos_prob_out,
                         ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_backward");
    if (!z || !lse || !workspace || !dz_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(z, 16) || !aligned(dz_out, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, M, opb, flags, tau);
    const DeviceWs w = carve_device((char*)workspace, M, D, opb, p);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, M, D, w.U, opb, w.inv_norm, stream, w.rows.counter,
                             (int)ticket_words(M)));
    return run_backward_device(z, dtype == NTXENT_BF16, w, opb, lse, grad_out, M, D, tau, p, dz_out,
                               pos_prob_out, stream);
}

int ntxent_b200_backward_tau(const void* z, const float* lse, const float* grad_out, int64_t M,
                             int64_t D, float tau, int dtype, unsigned flags, void* workspace,
                             size_t workspace_bytes, void* dz_out, float* tau_grad_rows_out,
                             ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_backward_tau");
    if (!z || !lse || !workspace || !dz_out || !tau_grad_rows_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(z, 16) || !aligned(dz_out, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, M, opb, flags, tau);
    const DeviceWs w = carve_device((char*)workspace, M, D, opb, p);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, M, D, w.U, opb, w.inv_norm, stream, w.rows.counter,
                             (int)ticket_words(M)));
    return run_backward_device(z, dtype == NTXENT_BF16, w, opb, lse, grad_out, M, D, tau, p, dz_out,
                               nullptr, stream, tau_grad_rows_out);
}

int ntxent_b200_forward_backward(const void* z, const float* grad_out, int64_t M, int64_t D,
                                 float tau, int dtype, unsigned flags, void* workspace,
                                 size_t workspace_bytes, float* loss_out, float* lse_out,
                                 void* dz_out, float* pos_prob_out, ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_forward_backward");
    if (!z || !workspace || !loss_out || !lse_out || !dz_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(z, 16) || !aligned(dz_out, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, M, opb, flags, tau);
    const DeviceWs w = carve_device((char*)workspace, M, D, opb, p);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, M, D, w.U, opb, w.inv_norm, stream, w.rows.counter,
                             (int)ticket_words(M)));
    rc = run_forward_device(w.U, opb, M, D, tau, p, w.rows, lse_out, loss_out, stream);
    if (rc != NTXENT_OK) return rc;
    return run_backward_device(z, dtype == NTXENT_BF16, w, opb, lse_out, grad_out, M, D, tau, p,
                               dz_out, pos_prob_out, stream);
}

int ntxent_b200_forward_backward_pos(const void* z, const int32_t* pos, const int32_t* inv_ptr,
                                     const int32_t* inv_idx, const float* grad_out, int64_t M, int64_t D,
                                     float tau, int dtype, unsigned flags, void* workspace,
                                     size_t workspace_bytes, float* loss_out, float* lse_out, void* dz_out,
                                     ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_forward_backward_pos");
    if (!z || !pos || !workspace || !loss_out || !lse_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (dz_out && (!inv_ptr || !inv_idx)) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(z, 16) || !aligned(dz_out, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, M, opb, flags, tau);
    const DeviceWs w = carve_device((char*)workspace, M, D, opb, p);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, M, D, w.U, opb, w.inv_norm, stream, w.rows.counter,
                             (int)ticket_words(M)));
    rc = run_forward_device(w.U, opb, M, D, tau, p, w.rows, lse_out, loss_out, stream, pos);
    if (rc != NTXENT_OK || !dz_out) return rc;
    return run_backward_device(z, dtype == NTXENT_BF16, w, opb, lse_out, grad_out, M, D, tau, p, dz_out, nullptr,
                               stream, nullptr, pos, inv_ptr, inv_idx);
}

int ntxent_b200_forward_rows(const void* u_all, int64_t M, int64_t D, int64_t row_begin,
                             int64_t row_count, float tau, int dtype, unsigned flags,
                             void* workspace, size_t workspace_bytes, float* lse_rows_out,
                             float* loss_partial_out, ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_forward_rows");
    if (!u_all || !workspace || !lse_rows_out || !loss_partial_out)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(u_all, 16) || !aligned(workspace, 256)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    if (row_begin < 0 || row_count < 1 || row_begin + row_count > M)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (flags & NTXENT_FLAG_CROSS_VIEW) return NTXENT_ERR_INVALID_ARGUMENT;  // whole-batch calls only
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, row_count, opb, flags, tau);
    const RowsWs w = carve_rows((char*)workspace, M, D, row_count, p, true);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    return run_forward_rows(u_all, opb, M, D, row_begin, row_count, tau, p, w, lse_rows_out,
                            loss_partial_out, 1.0f, stream);
}

int ntxent_b200_backward_rows(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                              int64_t row_begin, int64_t row_count, float tau, int mode, int dtype,
                              unsigned flags, void* workspace, size_t workspace_bytes,
                              float* du_rows_out, float* du_cols_out, ntxent_stream_t stream) {
    NTX_RANGE("ntxent_b200_backward_rows");
    if (!u_all || !lse_all || !workspace || !du_rows_out) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(u_all, 16) || !aligned(workspace, 256) || !aligned(du_rows_out, 16) ||
        !aligned(du_cols_out, 16))
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (mode != NTXENT_BWD_SYMMETRIC && mode != NTXENT_BWD_CONTRACT)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (mode == NTXENT_BWD_CONTRACT && !du_cols_out) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    if (row_begin < 0 || row_count < 1 || row_begin + row_count > M)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (flags & NTXENT_FLAG_CROSS_VIEW) return NTXENT_ERR_INVALID_ARGUMENT;  // whole-batch calls only
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, row_count, opb, flags, tau);
    const RowsWs w = carve_rows((char*)workspace, M, D, row_count, p, true);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    const bool sym = (mode == NTXENT_BWD_SYMMETRIC);
    rc = run_bwd_row_term(u_all, opb, lse_all, M, D, row_begin, row_count, tau, sym, p, w, du_rows_out, stream);
    if (rc != NTXENT_OK || sym) return rc;
    return run_bwd_col_term(u_all, opb, lse_all, M, D, row_begin, row_count, tau, p, w, du_cols_out, stream);
}

int ntxent_b200_backward_rows_part(const void* u_all, const float* lse_all, int64_t M, int64_t D,
                                   int64_t row_begin, int64_t row_count, float tau, int part,
                                   int dtype, unsigned flags, void* workspace, size_t workspace_bytes,
                                   float* out, ntxent_stream_t stream) {
    if (!u_all || !lse_all || !workspace || !out || (part != 0 && part != 1)) return NTXENT_ERR_INVALID_ARGUMENT;
    if (!aligned(u_all, 16) || !aligned(workspace, 256) || !aligned(out, 16)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    if (row_begin < 0 || row_count < 1 || row_begin + row_count > M)
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (flags & NTXENT_FLAG_CROSS_VIEW) return NTXENT_ERR_INVALID_ARGUMENT;  // whole-batch calls only
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, row_count, opb, flags, tau);
    const RowsWs w = carve_rows((char*)workspace, M, D, row_count, p, true);
    if (workspace_bytes < w.bytes) return NTXENT_ERR_WORKSPACE_TOO_SMALL;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    if (part == 0)
        return run_bwd_row_term(u_all, opb, lse_all, M, D, row_begin, row_count, tau, false, p, w, out, stream);
    return run_bwd_col_term(u_all, opb, lse_all, M, D, row_begin, row_count, tau, p, w, out, stream);
}

int ntxent_b200_finalize_dz(const float* du_rows, const void* z_rows, const float* inv_norm_rows,
                            const float* grad_out, int64_t rows, int64_t D, int dtype, void* dz_out,
                            ntxent_stream_t stream) {
    if (!du_rows || !z_rows || !inv_norm_rows || !dz_out || rows < 0 || !aligned(du_rows, 16) ||
        !aligned(z_rows, 16) || !aligned(dz_out, 16))
        return NTXENT_ERR_INVALID_ARGUMENT;
    if (dtype != NTXENT_F32 && dtype != NTXENT_BF16) return NTXENT_ERR_INVALID_ARGUMENT;
    if (D < 8 || (D % 8) != 0 || D > 1024) return NTXENT_ERR_UNSUPPORTED_SHAPE;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    NTX_TRY(launch_finalize_dz(du_rows, z_rows, dtype == NTXENT_BF16, inv_norm_rows, grad_out, rows,
                               D, dz_out, stream));
    return NTXENT_OK;
}

}  // extern "C"

// ---- forward for callers that cannot allocate --------------------------------------
// Everything below is an addition beside the entry points above; it shares their kernels and
// changes none of their code paths.
namespace {

// The scratch of the forward pass (normalised operands, 1/|z|, split partials, positive logits
// and the loss reduction's per-CTA partials) for ntxent_b200_forward_static.  It is part of
// the library image: every GPU whose context loads this library holds these bytes until the
// process exits, whether or not the entry point is ever called.
static __device__ __align__(256) unsigned char g_static_scratch[NTXENT_STATIC_SCRATCH_BYTES];

constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxStaticDevices] = {};
void* g_static_ctx[kMaxStaticDevices] = {};  // the driver context each address was looked up in

// The context that is current on this thread (nullptr when the driver call cannot be had).
void* current_context() {
    typedef int (*GetCurrentFn)(void**);
    static GetCurrentFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPointByVersion("cuCtxGetCurrent", &p, 12000, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            (void)cudaGetLastError();
            p = nullptr;
        }
        return reinterpret_cast<GetCurrentFn>(p);
    }();
    void* ctx = nullptr;
    if (!fn || fn(&ctx) != 0) return nullptr;
    return ctx;
}

// In host code `g_static_scratch` is only a handle for the runtime, not an address a kernel
// can use, so the device address is looked up once per device, with that device current.  The
// address belongs to the context it was looked up in: after cudaDeviceReset(), or when the host
// makes another context current on the same device, the remembered one is stale, so it is kept
// together with its context and looked up again when that changes (and on every call when the
// current context cannot be told).
int static_scratch_base(char** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxStaticDevices) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    std::lock_guard<std::mutex> lock(g_static_mu);
    void* ctx = current_context();
    if (!g_static_ptr[dev] || !ctx || ctx != g_static_ctx[dev]) {
        void* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_static_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ptr) return NTXENT_ERR_UNSUPPORTED_DEVICE;
        if (!ctx) ctx = current_context();  // the lookup above may have been what bound a context
        g_static_ptr[dev] = ptr;
        g_static_ctx[dev] = ctx;
    }
    *out = (char*)g_static_ptr[dev];
    return NTXENT_OK;
}

// Bytes of the workspace layout that the forward pass touches: everything in front of the
// backward accumulators.
size_t forward_scratch_bytes(int64_t M, int64_t D, bool op_bf16, const Plans& p) {
    char* const probe = reinterpret_cast<char*>(kAlign);  // offsets only, never dereferenced
    const DeviceWs w = carve_device(probe, M, D, op_bf16, p);
    return (size_t)(reinterpret_cast<char*>(w.rows.acc_rows) - probe);
}

}  // namespace

extern "C" int ntxent_b200_forward_static(const void* z, int64_t M, int64_t D, float tau, int dtype,
                                          unsigned flags, float* loss_out, float* lse_out,
                                          ntxent_stream_t stream) {
    if (!z || !loss_out || !lse_out || !aligned(z, 16)) return NTXENT_ERR_INVALID_ARGUMENT;
    int rc = check_shape(M, D, tau, dtype);
    if (rc != NTXENT_OK) return rc;
    const bool opb = op_is_bf16(dtype, flags);
    const Plans p = make_plans(M, D, M, opb, flags | kFlagNoSymForward, tau);
    if (forward_scratch_bytes(M, D, opb, p) > (size_t)NTXENT_STATIC_SCRATCH_BYTES)
        return NTXENT_ERR_STATIC_SCRATCH_EXCEEDED;
    if (check_device() != NTXENT_OK) return NTXENT_ERR_UNSUPPORTED_DEVICE;
    char* base = nullptr;
    rc = static_scratch_base(&base);
    if (rc != NTXENT_OK) return rc;
    const DeviceWs w = carve_device(base, M, D, opb, p);
    // Nothing in the array has to start at zero: every region is written before it is read, so
    // whatever the previous call left behind is harmless.
    NTX_TRY(launch_normalize(z, dtype == NTXENT_BF16, M, D, w.U, opb, w.inv_norm, stream, w.rows.counter,
                             (int)ticket_words(M)));
    return run_forward_rows(w.U, opb, M, D, 0, M, tau, p, w.rows, lse_out, loss_out,
                            1.0f / (float)M, stream);
}
