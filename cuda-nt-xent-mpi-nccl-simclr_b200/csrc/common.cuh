// common.cuh -- shared declarations of the ntxent_b200 library (no torch, no CUTLASS).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ntxent_b200 {

constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kNormEps = 1e-12f;  // torch.nn.functional.normalize's default eps

// Which weights the backward tile pass forms for a row a (of the A range) and a
// column b (of the B range); S_ab = u_a.u_b/tau, diagonal a == b always masked.
//   kSym : W_ab = exp(S_ab - l_a) + exp(S_ab - l_b)   (G + G^T, single device / symmetric form)
//   kRow : W_ab = exp(S_ab - l_a)                     (row term G_{I,:} U)
//   kCol : W_ab = exp(S_ab - l_b)                     (column term G_{I,:}^T U_I, a over all rows)
enum BwdMode : int { kSym = 0, kRow = 1, kCol = 2 };

// Output fan-out for the multi-GPU "push" path: the same values are stored to up to kMaxPeers
// buffers (this GPU's own and its peers' symmetric-memory copies, mapped over NVLink).
constexpr int kMaxPeers = 16;
struct PeerList {
    void* p[kMaxPeers];
    int n;
};
inline PeerList single_peer(void* ptr) {
    PeerList l{};
    l.p[0] = ptr;
    l.n = 1;
    return l;
}

// Cross-rank hand-off without a barrier kernel (multi-GPU push path).  Every rank owns an array
// of per-producer flags in symmetric memory.  A producing call stores its results into all the
// peers and then publishes its new epoch (a per-rank call counter kept in device memory, so that
// a replayed CUDA graph needs no new arguments) to slot `rank` of every peer's array -- from the
// one-block kernel that ends the call, i.e. after the grid that did the stores has completed.  A consuming kernel waits until the slots of the ranks whose rows it is about to
// read have reached the consumer's own epoch (all ranks make the same sequence of calls).
struct PeerSignal {
    PeerList flags;         // uint32 flag array [kMaxPeers] on every rank (this rank's own included)
    unsigned* epoch;        // local call counter; nullptr: no signalling
    int rank;
    unsigned* ticket;       // optional ticket counter for the fused forward finalize
};
struct PeerWait {
    const unsigned* flags;  // this rank's flag array; nullptr: no waiting
    const unsigned* epoch;  // local counter holding the value the slots must reach
    int ranks;
    int64_t rows_per_rank;  // rank r produced rows [r * rows_per_rank, (r + 1) * rows_per_rank)
    // A peer that stalls on its host (a checkpoint write, a debugger) must not kill this rank:
    // the wait gives up only after `timeout_ns` of %globaltimer (0: never), then sets *status to
    // 1 + the rank it was waiting for and carries on with whatever is in the buffer -- the host
    // reads the word (ntxent_b200_peer_wait_status) and knows the step's results are void.
    // Without a status word the kernel traps instead (the old behaviour, now after the timeout).
    unsigned long long timeout_ns;
    unsigned* status;
};
inline PeerSignal no_signal() { PeerSignal s{}; return s; }
inline PeerWait no_wait() { PeerWait w{}; return w; }

struct SplitPlan {
    int row_tiles;       // tiles over the A range
    int nsplit;          // column splits (grid.y)
    int64_t cols_per_split;  // multiple of the column tile
    // tcgen05 backward only: the tiles are dealt to `sk_ctas` persistent CTAs in equal contiguous
    // shares instead of a (row tile, split) grid; nsplit is then the maximum number of partials
    // of a row tile and sk_col_tiles the column-tile count the dealing was computed with.
    int sk_ctas;       // 0: plain grid split
    int sk_col_tiles;
};

// Column-split heuristic shared by the workspace sizing and the launchers: pick the
// split count whose last wave wastes the least, given `num_sms` concurrently running CTAs.
SplitPlan plan_splits(int64_t a_count, int64_t b_count, int row_tile, int col_tile, int num_sms,
                      int max_split);

int device_sm_count();  // SM count of the current device; 148 when there is none

// ---- launchers (all asynchronous on `stream`; return the launch status) ----------
// zero_words (optional): n_zero device words the kernel sets to zero -- the tickets of the forward
// finalize, cleared by the kernel that starts every whole-batch call.
cudaError_t launch_normalize(const void* z, bool z_bf16, int64_t rows, int64_t D, void* u_out,
                             bool op_bf16, float* inv_norm, cudaStream_t stream, unsigned* zero_words = nullptr,
                             int n_zero = 0);
// Same, storing the rows at row `row_offset` of every buffer in `outs`.
cudaError_t launch_normalize_push(const void* z, bool z_bf16, int64_t rows, int64_t D,
                                  const PeerList& outs, int64_t row_offset, bool op_bf16,
                                  float* inv_norm, const PeerSignal& signal, cudaStream_t stream,
                                  unsigned* zero_words = nullptr, int n_zero = 0);

// SIMT (fp32 FMA) tile kernels.
cudaError_t launch_simt_fwd(const void* u_all, bool op_bf16, int64_t M, int64_t D,
                            int64_t row_begin, int64_t row_count, float tau, const SplitPlan& plan,
                            float* part_m, float* part_s, const PeerWait& u_ready,
                            cudaStream_t stream, int64_t cross_n = 0);
cudaError_t launch_simt_bwd(const void* u_all, bool op_bf16, const float* lse_all, int64_t M,
                            int64_t D, int64_t a_begin, int64_t a_count, int64_t b_begin,
                            int64_t b_count, float tau, int mode, const SplitPlan& plan, float* acc,
                            const PeerWait& lse_ready, cudaStream_t stream, int64_t cross_n = 0);
// One warp that waits for every slot (for consumers that are not ours, e.g. a sum of the loss parts).
cudaError_t launch_peer_wait(const PeerWait& wait, cudaStream_t stream);

// Forward finalize: the positive logits k u_i . u_pos(i) (pos_index == nullptr: pos(i) =
// (i + M/2) mod M) into pos_t [row_count], then the split partials combined into l_i, the positive
// logit subtracted, the loss reduced.  l_i goes to element lse_offset + i of every buffer in
// `lse_outs`, the (scaled) loss to element loss_index of every buffer in `loss_outs` (one buffer
// each on a single device).  `counter`: a device word that is zero before the first call (and
// again after every call): with it the three kernels of this step run as one (a last-block ticket
// replaces the kernel boundaries); without it they are launched one after the other.
cudaError_t launch_fwd_finalize_push(const void* u_all, bool op_bf16, int64_t M, int64_t D,
                                     int64_t row_begin, int64_t row_count, float tau, int nsplit,
                                     const float* part_m, const float* part_s, float* pos_t, const int* pos_index,
                                     const PeerList& lse_outs, int64_t lse_offset, float* block_partials,
                                     const PeerList& loss_outs, int loss_index, float loss_scale,
                                     const PeerSignal& signal, cudaStream_t stream, unsigned* counter = nullptr);

// The loss: n per-CTA partials summed by one CTA in a fixed order, scaled, stored to element
// loss_index of every buffer in `loss_outs` (and, multi-GPU push path, published with `signal`).
cudaError_t launch_loss_reduce(const float* block_partials, int n, const PeerList& loss_outs, int loss_index,
                               float loss_scale, const PeerSignal& signal, cudaStream_t stream);

// pos_t[r] = (log2(e)/tau) u_i . u_pos(i), i = row_begin + r, as an fp32 dot product of the operand
// rows; pos(i) = pos_index[i], or (i + M/2) mod M when pos_index is null.
cudaError_t launch_pos_dot(const void* u_all, bool op_bf16, int64_t M, int64_t D, int64_t row_begin,
                           int64_t row_count, const int* pos_index, float k_scale, float* pos_t,
                           cudaStream_t stream);

struct BwdEpilogueArgs {
    const float* acc;        // [nsplit][a_count][D] raw sums  sum_b W_ab u_b
    int nsplit;
    int sk_ctas, sk_col_tiles;  // from the SplitPlan: per-row-tile partial counts when sk_ctas > 0
    const void* u_all;       // [M, D] operands
    bool op_bf16;
    int64_t M, D, a_begin, a_count;
    float pos_coeff;         // multiple of u_pos(a) to subtract (2 sym, 1 row/col)
    int64_t pos_lo, pos_hi;  // subtract only when pos(a) lies in [pos_lo, pos_hi)
    float scale;             // 1 / (tau * M)
    // output A: plain dU rows
    float* du_out;           // [a_count, D] or nullptr
    // output B: Jacobian applied (du_out == nullptr)
    const void* z_rows;      // [a_count, D] in z dtype
    bool z_bf16;
    const float* inv_norm;   // [a_count]
    const float* grad_out;   // device scalar or nullptr
    void* dz_out;            // [a_count, D] in z dtype
    // optional P_{a,pos(a)}
    const float* lse_all;    // [M] (needed only with pos_prob_out)
    float inv_tau;
    float* pos_prob_out;     // [a_count] or nullptr
    // optional: positives from an index vector instead of (a + M/2) mod M.  pos_index[a] is row a's
    // positive; rows b with pos_index[b] == a are inv_idx[inv_ptr[a] .. inv_ptr[a+1]).  The term
    // subtracted from row a is then u_pos[a] + sum of those u_b (pos_coeff / pos_lo / pos_hi unused).
    const int* pos_index;
    const int* inv_ptr;
    const int* inv_idx;
    // optional: loss_out[0] = loss_scale * (loss_parts[0] + ... + loss_parts[loss_n - 1]), by one thread
    // of the grid -- the global loss of the push path, whose parts are all there once the backward
    // tile kernel has run (it waited for every rank's flag)
    const float* loss_parts;
    int loss_n;
    float loss_scale;
    float* loss_out;
    // optional per-row temperature gradient (output B only): -(dU_a . u_a) / (2 tau); its sum
    // over all rows, times grad_out, is d loss / d tau
    float* tau_grad_rows_out;  // [a_count] or nullptr
};
cudaError_t launch_bwd_epilogue(const BwdEpilogueArgs& args, cudaStream_t stream);

cudaError_t launch_finalize_dz(const float* du_rows, const void* z_rows, bool z_bf16,
                               const float* inv_norm, const float* grad_out, int64_t rows, int64_t D,
                               void* dz_out, cudaStream_t stream);

// ---- launching ------------------------------------------------------------------
// Every kernel of the library is launched with programmatic stream serialisation (PDL): its
// CTAs may become resident while the kernel before it on the stream is still draining, run their
// set-up (barrier init, TMEM allocation, tensor-map prefetch) and then block in pdl_wait() until
// that kernel has completed and its stores are visible.  Each kernel calls pdl_trigger() first
// thing -- "my dependents may be scheduled once all my CTAs are running" -- and pdl_wait() before
// its first access to global memory that another kernel of the stream reads or writes (inputs of
// the step included: the previous step's last kernel may still be reading what this one is about
// to overwrite).  NTXENT_NO_PDL=1 in the environment launches the plain way (for A/B runs).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#ifdef __CUDACC__
// Row i's positive in the stacked-views layout, (i + M/2) mod M, without the 64-bit division
// (a hundred instructions per call in kernels that are a few hundred long).
__device__ __forceinline__ int64_t pair_row(int64_t i, int64_t M) {
    const int64_t h = M / 2, j = i + h;
    return j >= M ? j - M : j;
}
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
#endif

// ---- element access helpers ---------------------------------------------------
template <typename T>
struct Elem;
template <>
struct Elem<float> {
    static __device__ __forceinline__ float4 ld4(const float* p) {
        return *reinterpret_cast<const float4*>(p);
    }
    static __device__ __forceinline__ void st4(float* p, float4 v) {
        *reinterpret_cast<float4*>(p) = v;
    }
};
template <>
struct Elem<__nv_bfloat16> {
    static __device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
        uint2 raw = *reinterpret_cast<const uint2*>(p);
        __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&raw.x);
        __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
        float2 a = __bfloat1622float2(lo), b = __bfloat1622float2(hi);
        return make_float4(a.x, a.y, b.x, b.y);
    }
    static __device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
        __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 raw;
        raw.x = *reinterpret_cast<uint32_t*>(&lo);
        raw.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(p) = raw;
    }
};

#ifdef __CUDACC__
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long peer_globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Spin until rank r's slot reaches `want`.  The bound is wall time, not a poll count, and it is
// the caller's (default: ten minutes, the order of a collective-library watchdog): see PeerWait.
// Returns false when it gave up.
__device__ __forceinline__ bool peer_wait_rank(const PeerWait& w, int r, unsigned want) {
    if ((int)(ld_acquire_sys(w.flags + r) - want) >= 0) return true;
    unsigned spins = 0;
    unsigned long long t0 = 0;
    while ((int)(ld_acquire_sys(w.flags + r) - want) < 0) {
        __nanosleep(40);
        if ((++spins & 0x3ffu) == 0 && w.timeout_ns) {  // look at the clock every ~1000 polls
            const unsigned long long now = peer_globaltimer_ns();
            if (!t0) t0 = now;
            if (now - t0 > w.timeout_ns) {
                if (!w.status) __trap();
                atomicCAS(w.status, 0u, 1u + (unsigned)r);
                return false;
            }
        }
    }
    return true;
}
// Ranks that produced rows [lo, hi): wait for those not yet seen ready (bit r of `ready_mask`).
__device__ __forceinline__ void peer_wait_rows(const PeerWait& w, int64_t lo, int64_t hi, unsigned want,
                                               unsigned& ready_mask) {
    if (!w.flags || hi <= lo || ready_mask == 0xffffffffu) return;
    const int r0 = (int)min((int64_t)w.ranks - 1, lo / w.rows_per_rank);
    const int r1 = (int)min((int64_t)w.ranks - 1, (hi - 1) / w.rows_per_rank);
    for (int r = r0; r <= r1; ++r)
        if (!((ready_mask >> r) & 1u)) {
            // after a timeout nothing is waited for any more: the step is void, let it drain
            ready_mask = peer_wait_rank(w, r, want) ? (ready_mask | (1u << r)) : 0xffffffffu;
        }
}
// Whole warp: one round trip that polls every rank's slot at once (lane r reads slot r); returns
// the mask of ranks already published and the value to reach in `want`.  The common case -- all
// there -- then costs a consumer CTA one L2 access instead of one per rank.
__device__ __forceinline__ unsigned peer_poll_all_warp(const PeerWait& w, unsigned& want) {
    want = 0u;
    if (!w.flags) return 0xffffffffu;
    const int lane = threadIdx.x & 31;
    want = *w.epoch;
    const bool ok = lane >= w.ranks || (int)(ld_acquire_sys(w.flags + lane) - want) >= 0;
    return __ballot_sync(0xffffffffu, ok);
}
// Block-wide: wait for every rank (kernels without a dedicated loader warp).
__device__ __forceinline__ void peer_wait_all_block(const PeerWait& w) {
    if (!w.flags) return;
    if ((int)threadIdx.x < w.ranks) peer_wait_rank(w, threadIdx.x, *w.epoch);
    __syncthreads();
}
#endif

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Merge two (max, sum) pairs of a base-2 logsumexp; safe for empty (-inf, 0) partials.
__device__ __forceinline__ void lse2_merge(float& m, float& s, float m2, float s2) {
    float mn = fmaxf(m, m2);
    if (mn == -INFINITY) { m = mn; s = 0.f; return; }
    s = s * exp2f(m - mn) + s2 * exp2f(m2 - mn);
    m = mn;
}

}  // namespace ntxent_b200
