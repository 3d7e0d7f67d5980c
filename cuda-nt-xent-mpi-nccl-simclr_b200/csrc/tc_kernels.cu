// tc_kernels.cu -- the tcgen05 / TMEM / TMA tile kernels (bf16 operands, fp32 accumulation).
//
// Forward  (K1): per CTA one 128-row tile of U against a range of 128-column tiles.
//   TMA (128B swizzle) stages U tiles in shared memory; one thread issues tcgen05.mma
//   S = U_a U_b^T into a double-buffered TMEM accumulator; four warps pull S rows out with
//   tcgen05.ld (one row per thread) and fold them into a base-2 running (max, sum).
// Backward (K2): same S recompute; the four warps turn S into the weights W (bf16), write
//   them back over the S columns with tcgen05.st, and a second tcgen05.mma (A = W from TMEM,
//   B = the same U_b tile read MN-major) accumulates  acc_a += W_ab u_b  in TMEM.
//
// These replace the reference's cublasSgemm + row_max/softmax/compute_loss kernels
// (src/ntxent_kernel.cu:8-134,162-199) and its backward SGEMM (src/ntxent_kernel.cu:217-236).
#include "tc_kernels.cuh"

#include <cudaTypedefs.h>

#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <mutex>
#include <type_traits>

#include "tc_ptx.cuh"
#include "tc_tile.cuh"

namespace ntxent_b200 {
using namespace ptx;
using namespace tc;

namespace {

constexpr int kThreads = 320;               // warps 0..7 softmax, warp 8 TMA, warp 9 MMA
// The issuing warps carry the highest warp ids: the scheduler favours higher ids among ready
// warps, and a starved MMA issuer (a dozen scalar instructions per tcgen05.mma) stalls everything.
constexpr int kFwdTmaWarp = 8, kFwdMmaWarp = 9;
constexpr int kSoftmaxThreads = 256;        // two warps per TMEM lane quadrant, 64 columns each
constexpr int64_t kTf32MinRows = 1024;      // fp32 operands take the tf32 tiles from this many rows on

// One 32-column chunk of an S row folded into the running base-2 (max, sum).  The values are
// already t = log2(e)/tau * u_a.u_b: the A operand was scaled by k = log2(e)/tau when it was
// staged, which saves one FFMA per logit.  The fixed-max form sums 2^t as is (t <= k <= 50, so
// M terms of at most 2^50 stay far inside fp32) and reports max = 0.
template <bool kOnline>
__device__ __forceinline__ void fwd_chunk(const uint32_t (&v)[32], bool special, int64_t col0,
                                          int64_t row, int64_t M, float& m_run, float& s_run) {
    if (!kOnline) {
        float p0 = 0.f, p1 = 0.f;
        if (!special) {
            // packed pairs: one FADD2 per two logits, half of the pairs through the polynomial
            f32x2 a0 = 0ull, a1 = 0ull;
#pragma unroll
            for (int e = 0; e < 32; e += 4) {
                a0 = add2(a0, ex2_pair<kFwdPolyPairs>(v[e], v[e + 1], e >> 1));
                a1 = add2(a1, ex2_pair<kFwdPolyPairs>(v[e + 2], v[e + 3], (e >> 1) + 1));
            }
            float lo, hi;
            unpack2(add2(a0, a1), lo, hi);
            p0 = lo;
            p1 = hi;
        } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) {
                const int64_t col = col0 + e;
                p0 += (col < M && col != row) ? ex2(__uint_as_float(v[e])) : 0.f;
            }
        }
        s_run += p0 + p1;
    } else {
        float t[32], cmax = -INFINITY;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
            const int64_t col = col0 + e;
            const bool ok = !special || (col < M && col != row);
            t[e] = ok ? __uint_as_float(v[e]) : -INFINITY;
            cmax = fmaxf(cmax, t[e]);
        }
        const float mn = fmaxf(m_run, cmax);
        if (mn != -INFINITY) {
            float part = 0.f;
#pragma unroll
            for (int e = 0; e < 32; ++e) part += ex2(t[e] - mn);
            s_run = s_run * ex2(m_run - mn) + part;
            m_run = mn;
        }
    }
}

// ------------------------------------------------------------------ forward ---
template <int D, bool TF32, bool kOnline>
__global__ void __launch_bounds__(kThreads, Cfg<D, TF32>::kFwdCtasPerSm)
tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap, int64_t M, int64_t row_begin,
              int64_t row_count, float k_scale, int tiles_per_split, int col_tiles, int cross_half,
              float* __restrict__ part_m, float* __restrict__ part_s, PeerWait u_ready,
              const void* __restrict__ U_raw, FwdTail tail) {
    using C = Cfg<D, TF32>;
    constexpr int NS = C::kFwdStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    uint8_t* q_tile = smem;
    uint8_t* k_tiles = smem + C::kTileBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (1 + NS) * C::kTileBytes);
    uint64_t* q_full = bars;
    uint64_t* k_full = bars + 1;
    uint64_t* k_empty = k_full + NS;
    uint64_t* s_full = k_empty + NS;
    uint64_t* s_empty = s_full + 2;
    uint64_t* q_ready = s_empty + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(q_ready + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_trigger();

    const int64_t a0 = row_begin + (int64_t)blockIdx.x * kTile;
    // Which column tiles this CTA (row tile x, split y of gridDim.y) walks, and in which order:
    //   single device     : the contiguous range [y * tiles_per_split, ...)
    //   peers push U       : every gridDim.y-th tile, counted from the first tile of THIS rank's own
    //                        rows (which need no waiting) on through the ranks behind it, wrapping
    //                        around: every CTA starts on rows that are there already and meets the
    //                        peers' rows as late as possible, one rank after the other.
    //   cross-view form   : (single device only) col_tiles = T/2 tiles, those of the OTHER view: the
    //                        contiguous range shifted by cross_half = T/2 for the rows of view 1
    const bool rotated = u_ready.flags != nullptr;
    const int rot = rotated ? (int)((row_begin / kTile) % col_tiles) : 0;
    const int jt_first = (int)blockIdx.y * tiles_per_split;
    const int jt_shift = (cross_half > 0 && a0 < (int64_t)cross_half * kTile) ? cross_half : 0;
    const int jt0 = rotated ? (rot + (int)blockIdx.y) % col_tiles : jt_first + jt_shift;
    const int jt_step = rotated ? (int)gridDim.y : 1;
    const int n_tiles = rotated ? (col_tiles - (int)blockIdx.y + jt_step - 1) / jt_step
                                : min(tiles_per_split, col_tiles - jt_first);
    auto next_tile = [&](int jt) {  // rotated: (jt + jt_step) mod col_tiles, jt_step <= col_tiles
        jt += jt_step;
        return (rotated && jt >= col_tiles) ? jt - col_tiles : jt;
    };

    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmap);
        mbar_init(q_full, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], 8); }
        mbar_init(q_ready, 8);
        fence_mbar_init();
    }
    if (warp == kFwdMmaWarp) tmem_alloc<256>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // set-up done; from here on the kernel reads what the normalise kernel wrote

    if (warp == kFwdTmaWarp) {
        // multi-GPU push path: a tile is loaded once the ranks that produced its rows have
        // published them -- the CTA starts on the columns that are there instead of waiting
        // behind a barrier for all of them
        unsigned want;
        unsigned ready_mask = peer_poll_all_warp(u_ready, want);
        if (lane == 0) {
            // the flags are read through the generic proxy, the rows they cover through the async
            // proxy (TMA): a cross-proxy fence after every acquire that let new rows in
            unsigned fenced_mask = ~ready_mask;
            auto wait_rows = [&](int64_t lo, int64_t hi) {
                peer_wait_rows(u_ready, lo, hi, want, ready_mask);
                if (u_ready.flags && fenced_mask != ready_mask) {
                    fence_proxy_async_global();
                    fenced_mask = ready_mask;
                }
            };
            wait_rows(a0, min(M, a0 + kTile));
            mbar_arrive_expect_tx(q_full, C::kTileBytes);
            for (int a = 0; a < C::kAtoms; ++a)
                tma_load_2d(q_tile + a * kAtomBytes, &tmap, q_full, a * C::kAtomCols, (int32_t)a0);
            for (int it = 0, jt = jt0; it < n_tiles; ++it, jt = next_tile(jt)) {
                const int st = it % NS;
                const uint32_t ph = (it / NS) & 1;
                const int64_t c0 = (int64_t)jt * kTile;
                wait_rows(c0, min(M, c0 + kTile));
                mbar_wait(&k_empty[st], ph ^ 1);
                mbar_arrive_expect_tx(&k_full[st], C::kTileBytes);
                for (int a = 0; a < C::kAtoms; ++a)
                    tma_load_2d(k_tiles + st * C::kTileBytes + a * kAtomBytes, &tmap, &k_full[st],
                                a * C::kAtomCols, jt * kTile);
            }
        }
        __syncwarp();
    } else if (warp == kFwdMmaWarp) {
        // the whole warp walks the loop (uniform control flow); one elected lane issues
        const bool leader = elect_one();
        constexpr uint32_t idesc = idesc_op<TF32>(kTile, kTile, false);
        const uint32_t q_addr = smem_u32(q_tile);
        mbar_wait(q_ready, 0);
        for (int it = 0; it < n_tiles; ++it) {
            const int st = it % NS, buf = it & 1;
            mbar_wait(&k_full[st], (it / NS) & 1);
            mbar_wait(&s_empty[buf], ((it >> 1) & 1) ^ 1);
            tc_fence_after();
            const uint32_t k_addr = smem_u32(k_tiles + st * C::kTileBytes);
            if (leader) {
#pragma unroll
                for (int kk = 0; kk < C::kKSteps; ++kk)
                    mma_ss_f16<TF32>(tmem_base + buf * kTile, desc_kmajor(q_addr, kk), desc_kmajor(k_addr, kk),
                                     idesc, kk > 0);
                tc_commit(&k_empty[st]);
                tc_commit(&s_full[buf]);
            }
            __syncwarp();
        }
    } else {
        // eight softmax warps: warp w owns TMEM lanes 32*(w%4).. (one row per thread) and the
        // 64 columns [64h, 64h+64) of every S tile, h = w/4; with two CTAs per SM that is four warps
        // per scheduler, enough to cover each other's TMEM loads and barrier checks
        const int quad = warp & 3;
        const int half = warp >> 2;
        const int row_in_tile = quad * 32 + lane;
        const int64_t row = a0 + row_in_tile;
        const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + half * 64;
        float m_run = kOnline ? -INFINITY : 0.f;
        float s_run = 0.f;
        // scale the row tile by k = log2(e)/tau where TMA put it (elementwise, so the swizzle is
        // irrelevant), then publish it to the tensor core's (async-proxy) view of shared memory
        {
            mbar_wait(q_full, 0);
            uint4* q4 = reinterpret_cast<uint4*>(q_tile);
            for (int i = threadIdx.x; i < C::kTileBytes / 16; i += kSoftmaxThreads) {
                uint4 w = q4[i];
                uint32_t* p = reinterpret_cast<uint32_t*>(&w);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (TF32) {
                        p[j] = to_tf32(__uint_as_float(p[j]) * k_scale);
                    } else {
                        const float lo = __uint_as_float(p[j] << 16) * k_scale;
                        const float hi = __uint_as_float(p[j] & 0xffff0000u) * k_scale;
                        p[j] = pack_bf16(lo, hi);
                    }
                }
                q4[i] = w;
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_ready);
        }
        // two 32-column chunks per tile, double-buffered in registers: the load of the next chunk
        // (or of the next tile's first chunk) is in flight while the current one is exponentiated
        uint32_t va[32], vb[32];
        mbar_wait(&s_full[0], 0);
        tc_fence_after();
        tmem_ld32(lane_addr, va);
        // 64-column half tiles that need the masked path, as 32-bit numbers: those overlapping
        // this row tile's own rows (the diagonal) and everything from the ragged end of the range
        const int hj_diag_lo = (int)(a0 / 64), hj_diag_hi = (int)((a0 + 63) / 64) + 1;
        const int hj_ragged = (int)(M / 64);
        for (int it = 0, jt = jt0; it < n_tiles; ++it, jt = next_tile(jt)) {
            const int buf = it & 1;
            const int hj = 2 * jt + half;
            const bool special = (hj >= hj_diag_lo && hj <= hj_diag_hi) || hj >= hj_ragged;
            const int64_t j0 = (int64_t)hj * 64;  // (used on the masked path only)
            const uint32_t taddr = lane_addr + buf * kTile;
            tmem_ld_wait();
            tmem_ld32(taddr + 32, vb);
            fwd_chunk<kOnline>(va, special, j0, row, M, m_run, s_run);
            tmem_ld_wait();
            // this warp's S columns have left TMEM: hand the buffer back, then prefetch
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[buf]);
            if (it + 1 < n_tiles) {
                mbar_wait(&s_full[buf ^ 1], ((it + 1) >> 1) & 1);
                tc_fence_after();
                tmem_ld32(lane_addr + (buf ^ 1) * kTile, va);
            }
            fwd_chunk<kOnline>(vb, special, j0 + 32, row, M, m_run, s_run);
        }
        // merge the two column halves of each row (upper half hands over through shared memory)
        float2* halves = reinterpret_cast<float2*>(q_tile);  // the row tile is no longer needed:
        named_bar_sync(1, kSoftmaxThreads);                  // every MMA that read it has completed
        if (half == 1) halves[row_in_tile] = make_float2(m_run, s_run);
        named_bar_sync(1, kSoftmaxThreads);
        if (half == 0) {
            const float2 o = halves[row_in_tile];
            lse2_merge(m_run, s_run, o.x, o.y);
        }
        const int64_t r = (int64_t)blockIdx.x * kTile + row_in_tile;
        if (half == 0 && r < row_count) {
            part_m[(int64_t)blockIdx.y * row_count + r] = m_run;
            part_s[(int64_t)blockIdx.y * row_count + r] = s_run;
        }
        if (tail.tickets) {
            // ---- finalize in place of a kernel of its own (FwdTail, tc_kernels.cuh) ----
            unsigned* tick_s = reinterpret_cast<unsigned*>(q_tile + 2048);
            float* wsum_s = reinterpret_cast<float*>(q_tile + 2048 + 16);
            __threadfence();  // this CTA's partials, before its ticket
            named_bar_sync(1, kSoftmaxThreads);
            if (threadIdx.x == 0) tick_s[0] = atomicAdd(&tail.tickets[1 + blockIdx.x], 1u);
            named_bar_sync(1, kSoftmaxThreads);
            if (tick_s[0] == gridDim.y - 1) {  // the last split of this row tile to finish
                __threadfence();
                float contrib = 0.f;
                if (half == 0 && r < row_count) {
                    float m = -INFINITY, s = 0.f;
                    for (int sp = 0; sp < (int)gridDim.y; ++sp)
                        lse2_merge(m, s, __ldcg(part_m + (int64_t)sp * row_count + r),
                                   __ldcg(part_s + (int64_t)sp * row_count + r));
                    // the positive logit: fp32 dot of the two operand rows (both in L2)
                    const int64_t pi = tail.pos_index ? (int64_t)tail.pos_index[row] : pair_row(row, M);
                    float dot = 0.f;
                    if (TF32) {
                        const float4* ua = reinterpret_cast<const float4*>(static_cast<const float*>(U_raw) + row * D);
                        const float4* ub = reinterpret_cast<const float4*>(static_cast<const float*>(U_raw) + pi * D);
#pragma unroll 4
                        for (int k = 0; k < D / 4; ++k) {
                            const float4 a = ua[k], b = ub[k];
                            dot += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
                        }
                    } else {
                        const uint4* ua = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(U_raw) + row * D);
                        const uint4* ub = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(U_raw) + pi * D);
#pragma unroll 4
                        for (int k = 0; k < D / 8; ++k) {
                            const uint4 a = ua[k], b = ub[k];
                            const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                dot += __uint_as_float(aw[j] << 16) * __uint_as_float(bw[j] << 16) +
                                       __uint_as_float(aw[j] & 0xffff0000u) * __uint_as_float(bw[j] & 0xffff0000u);
                        }
                    }
                    const float l2 = m + log2f(s);
                    const float l = l2 * kLn2;
                    for (int i = 0; i < tail.lse_outs.n; ++i)
                        static_cast<float*>(tail.lse_outs.p[i])[tail.lse_offset + r] = l;
                    contrib = (l2 - dot * k_scale) * kLn2;
                }
                contrib = warp_sum(contrib);
                if (lane == 0) wsum_s[warp] = contrib;
                __threadfence_system();  // l_i at the peers, before the ticket that may publish them
                named_bar_sync(1, kSoftmaxThreads);
                if (threadIdx.x == 0) {
                    // (warps 0..3 hold the rows; 4..7 contributed zero)
                    tail.block_partials[blockIdx.x] = (wsum_s[0] + wsum_s[1]) + (wsum_s[2] + wsum_s[3]);
                    tail.tickets[1 + blockIdx.x] = 0u;  // ready for the next call
                    __threadfence_system();
                    tick_s[1] = atomicAdd(&tail.tickets[0], 1u);
                }
                named_bar_sync(1, kSoftmaxThreads);
                if (tick_s[1] == gridDim.x - 1 && warp == 0) {  // the last row tile: the loss, the signal
                    __threadfence();
                    double acc = 0.0;
                    for (int b = lane; b < (int)gridDim.x; b += 32) acc += (double)__ldcg(tail.block_partials + b);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
                    unsigned e = tail.signal.epoch ? *tail.signal.epoch + 1u : 0u;
                    __syncwarp();
                    if (lane == 0) {
                        if (tail.signal.epoch) *tail.signal.epoch = e;
                        tail.tickets[0] = 0u;
                    }
                    if (lane < tail.loss_outs.n) {
                        static_cast<float*>(tail.loss_outs.p[lane])[tail.loss_index] =
                            (float)(acc * (double)tail.loss_scale);
                        if (tail.signal.epoch)
                            st_release_sys(static_cast<unsigned*>(tail.signal.flags.p[lane]) + tail.signal.rank, e);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kFwdMmaWarp) tmem_dealloc<256>(tmem_base);
}

// ----------------------------------------------------------------- backward ---
// MODE: kSym / kRow / kCol (common.cuh).  kTwoExp (kSym only): W = 2^(t-l_a) + 2^(t-l_b) instead
// of the one-exponential form 2^t (2^-l_a + 2^-l_b), used when k = log2(e)/tau is so large that
// 2^t alone could leave the fp32 range.  t = k u_a.u_b comes straight out of the first MMA: the
// A operand is scaled by k when it is parked in TMEM.
//
// 640 threads: warps 0..15 weight stage, warp 16 TMA producer, warp 17 issues S = U_a U_b^T,
// warp 18 issues acc += W U_b, warp 19 stages the per-column constants a block ahead.  One
// persistent CTA per SM walks an equal share of the tiles (BwdUnit below).  Per 128x128
// tile the weight stage needs 16384 ex2 (MUFU: 16/clk/SM -> 1024 clk) and the tensor pipe 2 x 512
// clk, so the two must overlap completely and the weight warps must not move in lock-step:
// warp w owns TMEM lanes 32*(w%4).. (one row per thread) and columns [32q, 32q+32) of every S
// tile, q = w/4 -- four warps per scheduler that drift apart by up to half a tile, so one warp's
// barrier checks and TMEM traffic hide under the others' exponentials.
//
// The row tile U_a never touches shared memory: each thread loads its own row from global
// memory once and parks it in TMEM as the A operand of S = U_a U_b^T; both MMAs read A from TMEM
// and B from the one shared-memory copy of U_b (K-major for S, MN-major for W U_b).
//
// TMEM columns, D <= 128:  S (128) | W0 (64) | W1 (64) | acc (D) | U_a (D/2)
//   S is handed back as soon as all weight warps have pulled their columns into registers, so
//   S(it+1) is computed while tile it is exponentiated; the bf16 weights alternate between two
//   blocks, so storing W(it) never waits for the second MMA of tile it-1.
// TMEM columns, D  > 128:  S/W (128, W overwrites the S columns it came from) | acc (D) | U_a (D/2)
//   no room for separate W blocks: S(it+1) has to wait for the second MMA of tile it.
// tf32 (fp32 operands): the weights stay 32-bit, so W always overwrites S in place (128 columns)
// and U_a takes D columns:  S/W (128) | acc (D) | U_a (D).
template <int D, bool TF32 = false>
struct BwdCfg {
    static constexpr bool kSeparateP = !TF32 && (D <= 128);
    // separate W blocks come with TWO S buffers (S(g+1) is issued while the weight warps are still
    // reading S(g)); W gets two blocks when the columns are there (D <= 64) and one otherwise
    static constexpr int kSBufs = kSeparateP ? 2 : 1;
    static constexpr int kPBufs = kSeparateP ? (D <= 64 ? 2 : 1) : 1;
    static constexpr uint32_t kPCol = 128 * kSBufs;
    static constexpr uint32_t kAccCol = kSeparateP ? kPCol + 64 * kPBufs : 128;
    static constexpr uint32_t kQCol = kAccCol + D;
    static constexpr uint32_t kQCols = TF32 ? D : D / 2;
    static_assert(kQCol + kQCols <= 512, "TMEM budget");
    // bf16: one ring of U_b tiles feeds both MMAs (the second one reads the tile MN-major).
    // tf32: 32-bit operands cannot be read transposed out of a 16-byte-swizzled tile, so the
    // second MMA gets its B operand K-major from a transposed copy U^T: a one-deep ring of U_b
    // tiles (free again right after the S MMA) plus a ring of U^T tiles (free after W U).
    static constexpr int kStages = TF32 ? 1 : ((Cfg<D, TF32>::kTileBytes <= 32 * 1024) ? 4 : 3);
    static constexpr int kTStages = TF32 ? 2 : 0;
    static constexpr int kTAtomBytes = D * 128;             // U^T atom: [D rows x 32 fp32]
    static constexpr int kTTileBytes = 4 * kTAtomBytes;     // [D rows x 128 columns of the tile]
    // column constants (l_b or 2^-l_b) are staged by their own warp, kCcTiles tiles at a time into
    // one of two blocks, a block ahead of the weight warps
    static constexpr int kCcTiles = 16;
    static constexpr int kSmem =
        kStages * Cfg<D, TF32>::kTileBytes + kTStages * kTTileBytes + 2048 + 2 * kCcTiles * kTile * 4;
    static_assert(kSmem <= 227 * 1024, "shared memory budget");
    // 16 weight warps, the TMA warp, then two MMA issuers: tcgen05.mma issue blocks while the
    // pipe is busy, so one thread issuing both S(it+1) and W(it) U would tie the two hand-off
    // chains together (S(it+1) could not go out until W(it-1) U had been fed in)
    static constexpr int kThreads = 640;
    static constexpr int kTmaWarp = 16, kMmaWarp = 17, kMmaPvWarp = 18, kConstWarp = 19;
    static constexpr int kWeightThreads = 512;
};

// kPacked: bf16 pairs into out[0..15]; otherwise 32 fp32/tf32 weights into out[0..31] (in place
// over the S values when out aliases v).
template <int MODE, bool kTwoExp, bool kSpecial, bool kPacked = true, int NOUT = 16>
__device__ __forceinline__ void weight_chunk(const uint32_t (&v)[32], uint32_t (&packed)[NOUT],
                                             const float* __restrict__ cc, float k_scale, float row_c,
                                             int64_t col0, int64_t row, int64_t b_end) {
    static_assert(NOUT == (kPacked ? 16 : 32), "output size");
#ifndef NTX_BWD_UNPACKED  // (A/B build switch)
    if (MODE == kSym && !kTwoExp && !kSpecial && kBwdPoly == 0) {
        // the common case, on packed pairs: 2^t (2^-l_a + 2^-l_b) as one FADD2 + one FMUL2 per two
        // weights instead of two FADD + two FMUL
        const f32x2 rc2 = pack2(row_c, row_c);
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
            const f32x2 sum = add2(rc2, pack2(cc[e], cc[e + 1]));
            const f32x2 x = mul2(ex2_pair<kBwdPolyPairs>(v[e], v[e + 1], e >> 1), sum);
            float w0, w1;
            unpack2(x, w0, w1);
            if (kPacked) {
                packed[e >> 1] = pack_bf16(w0, w1);
            } else {
                packed[e] = to_tf32(w0);
                packed[e + 1] = to_tf32(w1);
            }
        }
        return;
    }
#endif
#pragma unroll
    for (int e = 0; e < 32; e += 2) {
        float w[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float t = __uint_as_float(v[e + h]);  // k u_a.u_b (the A operand carries k)
            float x;
            if (MODE == kRow) {
                x = ex2(t - row_c);
            } else if (MODE == kCol) {
                x = ex2(t - cc[e + h]);
            } else if (kTwoExp) {
                x = ex2(t - row_c) + ex2(t - cc[e + h]);
            } else {
                x = ex2_mix<kBwdPoly>(t, e + h) * (row_c + cc[e + h]);  // 2^t (2^-l_a + 2^-l_b)
            }
            if (kSpecial) {
                const int64_t col = col0 + e + h;
                if (col >= b_end || col == row) x = 0.f;
            }
            w[h] = x;
        }
        if (kPacked) {
            packed[e >> 1] = pack_bf16(w[0], w[1]);
        } else {
            packed[e] = to_tf32(w[0]);
            packed[e + 1] = to_tf32(w[1]);
        }
    }
}

// Optional pipeline trace (tools/trace_bwd.py): CTA 0 stamps clock64() at its hand-offs into
// trace[role][tile][event] (5 x 64 x 8 words); role 0 = MMA warps, 1 = weight warp 0, 2 = weight
// warp 15, 3 = the other weight warps (overwriting each other), all for the CTA's first unit;
// role 4 = unit-level events of weight warp 0 (row 63: kernel entry / set-up done / exit, with
// %globaltimer in events 6 and 7 to calibrate the clock).
// Compiled in only with -DNTXENT_TRACE=1 (NTXENT_TRACE=1 in the environment of the build, as
// tools/trace_bwd.py does): even predicated off, the stamps cost issue slots in the weight loop.
constexpr int kTraceTiles = 64, kTraceEvents = 8;
unsigned long long* g_trace = nullptr;
#ifndef NTXENT_TRACE
#define NTXENT_TRACE 0
#endif
#if NTXENT_TRACE
#define NTX_TRACE(role, it, ev)                                                                   \
    do {                                                                                          \
        if (trace && blockIdx.x == 0 && un == 0 && (it) < kTraceTiles && lane == 0)               \
            trace[((role)*kTraceTiles + (it)) * kTraceEvents + (ev)] = clock64();                 \
    } while (0)

#define NTX_TRACE_U(row, ev, val)                                                                  \
    do {                                                                                          \
        if (trace && blockIdx.x == 0 && warp == 0 && lane == 0 && (row) < kTraceTiles)            \
            trace[(4 * kTraceTiles + (row)) * kTraceEvents + (ev)] = (val);                       \
    } while (0)
#else
#define NTX_TRACE(role, it, ev) do { (void)trace; } while (0)
#define NTX_TRACE_U(row, ev, val) do { (void)trace; } while (0)
#endif
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Work distribution of the backward tile pass.  The (row tile, column tile) pairs are numbered
// row-major, t = rt * col_tiles + ct, and CTA c of `ctas` takes the contiguous range
// [c*T/ctas, (c+1)*T/ctas): every CTA gets the same number of tiles to within one, whatever the
// shape.  A CTA's range is cut at row-tile boundaries into *units*; each unit accumulates its own
// partial of the row tile's  sum_b W_ab u_b  and stores it to slot  c - (first CTA touching that
// row tile)  of acc_out, so the epilogue can add the partials of a row in a fixed order.
struct BwdUnit {
    int rt;    // row tile
    int ct0;   // first column tile
    int n;     // number of column tiles
    int slot;  // partial slot of this row tile
};
__host__ __device__ __forceinline__ int64_t bwd_cta_of_tile(int64_t t, int64_t total, int ctas) {
    return ((t + 1) * ctas - 1) / total;
}
__device__ __forceinline__ bool bwd_next_unit(int64_t& t, int64_t t_end, int col_tiles, int64_t total,
                                              int ctas, BwdUnit& u) {
    if (t >= t_end) return false;
    u.rt = (int)(t / col_tiles);
    const int64_t first = (int64_t)u.rt * col_tiles;
    u.ct0 = (int)(t - first);
    u.n = (int)min((int64_t)(col_tiles - u.ct0), t_end - t);
    u.slot = (int)blockIdx.x - (int)bwd_cta_of_tile(first, total, ctas);
    t += u.n;
    return true;
}

template <int D, bool TF32, int MODE, bool kTwoExp>
__global__ void __launch_bounds__(BwdCfg<D, TF32>::kThreads, 1)
tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_t,
              const void* __restrict__ U_raw,
              const float* lse, int64_t M, int64_t a_begin, int64_t a_count,
              int64_t b_begin, int64_t b_count, float k_scale, int col_tiles, int64_t total_tiles,
              int cross_half, float* __restrict__ acc_out, unsigned long long* __restrict__ trace,
              PeerWait lse_ready, BwdTail tail) {
    using C = Cfg<D, TF32>;
    using B = BwdCfg<D, TF32>;
    constexpr int NS = B::kStages;
    constexpr bool SEP = B::kSeparateP;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = align1024(smem_raw);
    constexpr int NT = B::kTStages;
    uint8_t* k_tiles = smem;
    uint8_t* kt_tiles = smem + NS * C::kTileBytes;  // tf32 only: U^T tiles for the second MMA
    float* colc = reinterpret_cast<float*>(kt_tiles + NT * B::kTTileBytes);  // [2][kCcTiles][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(colc + 2 * B::kCcTiles * kTile);
    uint64_t* q_full = bars;
    uint64_t* k_full = bars + 1;
    uint64_t* k_empty = k_full + NS;
    uint64_t* kt_full = k_empty + NS;   // [2] (tf32)
    uint64_t* kt_empty = kt_full + 2;   // [2] (tf32)
    uint64_t* s_full = kt_empty + 2;    // [2] S tile complete (MMA commit)
    uint64_t* s_free = s_full + 2;     // [2] S buffer reusable: weight warps (SEP) or second MMA (aliased W)
    uint64_t* p_full = s_free + 2;     // [2] weights stored by all sixteen warps
    uint64_t* p_free = p_full + 2;     // [2] second MMA has consumed the W block (SEP only)
    uint64_t* acc_full = p_free + 2;   // a unit's accumulator is complete (commit after its last W U)
    uint64_t* acc_free = acc_full + 1; // ... and has been read out by all sixteen weight warps
    uint64_t* cc_full = acc_free + 1;  // [2] a block of column constants is staged
    uint64_t* cc_empty = cc_full + 2;  // [2] ... and all sixteen weight warps are done with it
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(cc_empty + 2);
    volatile uint32_t* tail_flag = tmem_slot + 1;  // in-kernel epilogue: "this CTA took the last ticket"

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    pdl_trigger();
    NTX_TRACE_U(63, 0, clock64());
    NTX_TRACE_U(63, 6, globaltimer_ns());
    const int64_t a_end = a_begin + a_count;
    // The B range of a unit.  Cross-view form (cross_half = T/2 > 0, b_count = M/2): the row tiles of
    // view 1 take the columns of view 2 and the other way round; otherwise [b_begin, b_begin + b_count).
    auto unit_b_begin = [&](int rt) -> int64_t {
        if (cross_half > 0 && a_begin / kTile + rt < cross_half) return b_begin + (int64_t)cross_half * kTile;
        return b_begin;
    };
    const int ctas = (int)gridDim.x;
    const int64_t t_begin = (int64_t)blockIdx.x * total_tiles / ctas;
    const int64_t t_end = (int64_t)(blockIdx.x + 1) * total_tiles / ctas;
    // Every role walks the same units; `g` counts tiles across units (ring slots and barrier
    // phases carry on from one unit to the next), `un` counts units.
    int64_t t_next = t_begin;
    BwdUnit u;
    int g = 0, un = 0;

    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmap);
        mbar_init(q_full, 16);
        for (int s = 0; s < NS; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { mbar_init(&kt_full[s], 1); mbar_init(&kt_empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_free[b], SEP ? 16 : 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&p_full[b], 16); mbar_init(&p_free[b], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_free, 16);
        for (int b = 0; b < 2; ++b) { mbar_init(&cc_full[b], 1); mbar_init(&cc_empty[b], 16); }
        fence_mbar_init();
    }
    if (warp == B::kMmaWarp) tmem_alloc<512>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // set-up done; U, l and the accumulator slots belong to earlier kernels until here
    NTX_TRACE_U(63, 1, clock64());

    if (warp == B::kTmaWarp) {
        if (lane == 0) {
            while (bwd_next_unit(t_next, t_end, col_tiles, total_tiles, ctas, u)) {
                const int64_t ub = unit_b_begin(u.rt);
                for (int it = 0; it < u.n; ++it, ++g) {
                    const int st = g % NS;
                    const int32_t col0 = (int32_t)(ub + (int64_t)(u.ct0 + it) * kTile);
                    mbar_wait(&k_empty[st], ((g / NS) & 1) ^ 1);
                    mbar_arrive_expect_tx(&k_full[st], C::kTileBytes);
                    for (int a = 0; a < C::kAtoms; ++a)
                        tma_load_2d(k_tiles + st * C::kTileBytes + a * kAtomBytes, &tmap, &k_full[st],
                                    a * C::kAtomCols, col0);
                    if (TF32) {  // the same columns of U^T: four [D rows x 32 columns] atoms
                        const int tt = g & 1;
                        mbar_wait(&kt_empty[tt], ((g >> 1) & 1) ^ 1);
                        mbar_arrive_expect_tx(&kt_full[tt], B::kTTileBytes);
                        for (int a = 0; a < 4; ++a)
                            tma_load_2d(kt_tiles + tt * B::kTTileBytes + a * B::kTAtomBytes, &tmap_t,
                                        &kt_full[tt], col0 + a * 32, 0);
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == B::kMmaWarp) {
        // S issuer.  The whole warp walks the loop (uniform control flow); one elected lane issues.
        const bool leader = elect_one();
        constexpr uint32_t idesc_s = idesc_op<TF32>(kTile, kTile, false);
        for (; bwd_next_unit(t_next, t_end, col_tiles, total_tiles, ctas, u); ++un) {
            mbar_wait(q_full, un & 1);  // this unit's k U_a is parked in TMEM
            tc_fence_after();
            for (int it = 0; it < u.n; ++it, ++g) {
                const int st = g % NS;
                NTX_TRACE(0, it, 0);
                mbar_wait(&k_full[st], (g / NS) & 1);
                // separate W blocks: S buffer g % 2 is free once the weight warps have read S(g-2);
                // aliased W: the one buffer is free once the second MMA of tile g-1 has consumed
                // the weights stored over it
                const int sb = g % B::kSBufs;
                mbar_wait(&s_free[sb], ((g / B::kSBufs) & 1) ^ 1);
                tc_fence_after();
                NTX_TRACE(0, it, 2);
                const uint32_t k_addr = smem_u32(k_tiles + st * C::kTileBytes);
                if (leader) {
#pragma unroll
                    for (int kk = 0; kk < C::kKSteps; ++kk)
                        mma_ts_f16<TF32>(tmem_base + sb * 128, tmem_base + B::kQCol + kk * 8,
                                         desc_kmajor(k_addr, kk), idesc_s, kk > 0);
                    if (TF32) tc_commit(&k_empty[st]);  // tf32: only this MMA reads the U_b tile
                    tc_commit(&s_full[sb]);
                }
                __syncwarp();
                NTX_TRACE(0, it, 3);
            }
        }
    } else if (warp == B::kMmaPvWarp) {
        // acc += W(it) U_b issuer
        const bool leader = elect_one();
        constexpr uint32_t idesc_pv = idesc_op<TF32>(kTile, D, /*b_mn_major=*/!TF32);
        constexpr int kPvSteps = kTile / C::kKPerMma;  // 8 (bf16) or 16 (tf32) MMAs over the tile's 128 K rows
        for (; bwd_next_unit(t_next, t_end, col_tiles, total_tiles, ctas, u); ++un) {
        if (un > 0) {  // the previous unit's accumulator has been read out
            mbar_wait(acc_free, (un - 1) & 1);
            tc_fence_after();
        }
        for (int it = 0; it < u.n; ++it, ++g) {
            const int st = g % NS;
            const int pb = g % B::kPBufs;
            NTX_TRACE(0, it, 4);
            mbar_wait(&p_full[pb], (g / B::kPBufs) & 1);
            tc_fence_after();
            NTX_TRACE(0, it, 5);
            if (TF32) mbar_wait(&kt_full[g & 1], (g >> 1) & 1);
            const uint32_t k_addr = TF32 ? smem_u32(kt_tiles + (g & 1) * B::kTTileBytes)
                                         : smem_u32(k_tiles + st * C::kTileBytes);
            if (leader) {
#pragma unroll
                for (int kk = 0; kk < kPvSteps; ++kk) {
                    // W for K rows [32q, 32q+32): separate -> columns [16q, 16q+16) of block pb;
                    // aliased bf16 -> columns [32q, 32q+16) of the S buffer; tf32 -> column = K row
                    const uint32_t a_col = TF32 ? kk * 8
                                                : (SEP ? B::kPCol + pb * 64 + kk * 8
                                                       : (kk >> 1) * 32 + (kk & 1) * 8);
                    // B: bf16 -> the U_b tile read MN-major; tf32 -> the U^T tile, K-major (rows = d,
                    // 32-byte slices of 8 tile columns, atoms of D rows)
                    const uint64_t b_desc =
                        TF32 ? smem_desc_sw128(k_addr + (kk >> 2) * B::kTAtomBytes + (kk & 3) * 32, 16, 1024)
                             : desc_mnmajor<false>(k_addr, kk);
                    mma_ts_f16<TF32>(tmem_base + B::kAccCol, tmem_base + a_col, b_desc, idesc_pv,
                                     (it > 0) || (kk > 0));
                }
                tc_commit(TF32 ? &kt_empty[g & 1] : &k_empty[st]);
                tc_commit(SEP ? &p_free[pb] : &s_free[0]);
            }
            __syncwarp();
            NTX_TRACE(0, it, 6);
        }
        if (leader) tc_commit(acc_full);
        __syncwarp();
        }
    } else if (warp == B::kConstWarp) {
        // column constants (l_b in base-2 units, or e_b = 2^-l_b), one block of kCcTiles tiles at a
        // time; `cb` counts blocks across units
        if (MODE != kRow) {
            int cb = 0;
            // multi-GPU push path: a block of l_b is read once the ranks that produced it have
            // published it (the rows of the A range are this rank's own and need no wait)
            unsigned want;
            unsigned ready_mask = peer_poll_all_warp(lse_ready, want);
            while (bwd_next_unit(t_next, t_end, col_tiles, total_tiles, ctas, u)) {
                const int64_t ub = unit_b_begin(u.rt), b_end = ub + b_count;
                for (int it0 = 0; it0 < u.n; it0 += B::kCcTiles, ++cb) {
                    float* dst = colc + (cb & 1) * B::kCcTiles * kTile;
                    const int ncols = min(B::kCcTiles, u.n - it0) * kTile;
                    const int64_t c0 = ub + (int64_t)(u.ct0 + it0) * kTile;
                    peer_wait_rows(lse_ready, c0, min(b_end, c0 + ncols), want, ready_mask);
                    mbar_wait(&cc_empty[cb & 1], ((cb >> 1) & 1) ^ 1);
                    // four columns per load, eight loads in flight per lane: two round trips per
                    // full block (scalar loads when the range start is not 16-byte aligned)
                    auto cval = [&](float l) {
                        const float l2 = l * kLog2e;  // past the range (INFINITY): W = 0 either way
                        return (MODE == kSym && !kTwoExp) ? ex2(-l2) : l2;
                    };
                    const bool vec_ok = (c0 & 3) == 0;
                    for (int base = 0; base < ncols; base += 8 * 128) {
                        float4 l[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int c = base + j * 128 + lane * 4;
                            l[j] = make_float4(INFINITY, INFINITY, INFINITY, INFINITY);
                            if (c < ncols) {
                                // (peers may still be storing other blocks of l while this kernel
                                // runs: coherent loads, never ld.global.nc)
                                if (vec_ok && c0 + c + 3 < b_end) {
                                    l[j] = ld_f4_coherent(lse + c0 + c);
                                } else {
                                    if (c0 + c < b_end) l[j].x = ld_f1_coherent(lse + c0 + c);
                                    if (c0 + c + 1 < b_end) l[j].y = ld_f1_coherent(lse + c0 + c + 1);
                                    if (c0 + c + 2 < b_end) l[j].z = ld_f1_coherent(lse + c0 + c + 2);
                                    if (c0 + c + 3 < b_end) l[j].w = ld_f1_coherent(lse + c0 + c + 3);
                                }
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int c = base + j * 128 + lane * 4;
                            if (c < ncols)
                                *reinterpret_cast<float4*>(dst + c) =
                                    make_float4(cval(l[j].x), cval(l[j].y), cval(l[j].z), cval(l[j].w));
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&cc_full[cb & 1]);
                }
            }
            // in-kernel epilogue, multi-GPU form: the loss partials travel with l, so once every
            // rank's flag has been seen CTA 0 adds them in rank order
            if (tail.loss_out && blockIdx.x == 0) {
                peer_wait_rows(lse_ready, 0, M, want, ready_mask);
                if (lane == 0) {
                    float acc = 0.f;
                    for (int i = 0; i < tail.loss_n; ++i) acc += ld_f1_coherent(tail.loss_parts + i);
                    *tail.loss_out = acc * tail.loss_scale;
                }
            }
        }
    } else {
        const int quad = warp & 3;
        const int quarter = warp >> 2;  // which 32 columns of every S tile
        const int row_in_tile = quad * 32 + lane;
        const int wt = threadIdx.x;  // 0..511: the weight threads come first
        const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16);
        const int trole = warp == 0 ? 1 : (warp == 8 ? 2 : 3);  // two warps of scheduler 0
        (void)trole;

        // park k U_a of row tile `rt` in TMEM (bf16 pairs, or fp32 for tf32): column groups of the
        // row dealt round-robin to the four warps that share this lane quadrant.  Called for the
        // next unit once this warp has seen the last S tile of the current one complete, i.e. when
        // no MMA reads the old rows any more.
        auto park_q = [&](int rt) {
            const int64_t row = a_begin + (int64_t)rt * kTile + row_in_tile;
            if (TF32) {
                const float4* src = reinterpret_cast<const float4*>(static_cast<const float*>(U_raw) + row * D);
                for (int gq = quarter; gq < D / 32; gq += 4) {  // 32 columns = 32 floats = 8 x float4
                    uint32_t p[32];
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (row < M) t = __ldg(src + gq * 8 + q);
                        p[q * 4 + 0] = to_tf32(t.x * k_scale);
                        p[q * 4 + 1] = to_tf32(t.y * k_scale);
                        p[q * 4 + 2] = to_tf32(t.z * k_scale);
                        p[q * 4 + 3] = to_tf32(t.w * k_scale);
                    }
                    tmem_st32(lane_addr + B::kQCol + gq * 32, p);
                }
            } else {
                const uint4* src =
                    reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(U_raw) + row * D);
                for (int gq = quarter; gq < D / 32; gq += 4) {  // 16 packed columns = 32 bf16 = 4 x uint4
                    uint32_t p[16];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uint4 t = make_uint4(0u, 0u, 0u, 0u);
                        if (row < M) t = __ldg(src + gq * 4 + q);
                        const uint32_t raw[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
                        for (int j = 0; j < 4; ++j)  // scaled by k = log2(e)/tau: S comes out as k u_a.u_b
                            p[q * 4 + j] = pack_bf16(__uint_as_float(raw[j] << 16) * k_scale,
                                                     __uint_as_float(raw[j] & 0xffff0000u) * k_scale);
                    }
                    tmem_st16(lane_addr + B::kQCol + gq * 16, p);
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(q_full);
        };

        // The epilogue of a row tile, by the CTA whose unit completes it (BwdTail, tc_kernels.cuh).
        // Called by all sixteen weight warps after the unit's partial has been stored: partials in
        // slot order, the positive-pair term, the Jacobian of the normalisation, dz -- the arithmetic
        // of bwd_epilogue_kernel (epilogue_kernels.cu), row for row.
        auto tail_unit = [&](const BwdUnit& uu) {
            const int64_t first = (int64_t)uu.rt * col_tiles;
            const int nslots = (int)(bwd_cta_of_tile(first + col_tiles - 1, total_tiles, ctas) -
                                     bwd_cta_of_tile(first, total_tiles, ctas)) + 1;
            if (nslots > 1) {
                __threadfence();  // this unit's partial, before its ticket
                named_bar_sync(2, B::kWeightThreads);
                if (wt == 0) {
                    const bool last = atomicAdd(&tail.tickets[uu.rt], 1u) == (unsigned)(nslots - 1);
                    if (last) tail.tickets[uu.rt] = 0u;  // left zero for the next launch
                    *tail_flag = last ? 1u : 0u;
                }
                named_bar_sync(2, B::kWeightThreads);
                if (*tail_flag == 0u) return;
                __threadfence();  // the other units' partials, after their tickets
            } else {
                // the only partial is this CTA's own, but it was stored by other threads than the
                // ones that read it below (thread-per-row there, warp-per-rows here)
                named_bar_sync(2, B::kWeightThreads);
            }
            // warp w takes rows [8w, 8w + 8) of the tile, four at a time; a lane holds four columns
            // of each (coalesced 512-byte rows); every load of a batch goes out before the first use
            const int k = lane * 4;
            const float g = tail.grad_out ? *tail.grad_out : 1.0f;
            const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
            for (int batch = 0; batch < 2; ++batch) {
                const int64_t rb = (int64_t)uu.rt * kTile + warp * 8 + batch * 4;
                if (rb >= a_count) break;
                float4 t0[4], t1[4], up[4], zz[4];
                float inv[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int64_t r = min(rb + q, a_count - 1);  // rows past the end redo the last row, unstored
                    t0[q] = t1[q] = up[q] = zz[q] = zero4;
                    inv[q] = tail.inv_norm[r];
                    if (k < D) {
                        const float* src = acc_out + r * D + k;
                        t0[q] = __ldcg(reinterpret_cast<const float4*>(src));
                        if (nslots > 1) t1[q] = __ldcg(reinterpret_cast<const float4*>(src + a_count * D));
                        const uint2 ur = __ldg(reinterpret_cast<const uint2*>(
                            static_cast<const __nv_bfloat16*>(U_raw) + pair_row(a_begin + r, M) * D + k));
                        up[q] = make_float4(__uint_as_float(ur.x << 16), __uint_as_float(ur.x & 0xffff0000u),
                                            __uint_as_float(ur.y << 16), __uint_as_float(ur.y & 0xffff0000u));
                        if (tail.z_bf16) {
                            const uint2 zr = __ldg(reinterpret_cast<const uint2*>(
                                static_cast<const __nv_bfloat16*>(tail.z_rows) + r * D + k));
                            zz[q] = make_float4(__uint_as_float(zr.x << 16), __uint_as_float(zr.x & 0xffff0000u),
                                                __uint_as_float(zr.y << 16), __uint_as_float(zr.y & 0xffff0000u));
                        } else {
                            zz[q] = __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(tail.z_rows) + r * D + k));
                        }
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    t0[q].x += t1[q].x; t0[q].y += t1[q].y; t0[q].z += t1[q].z; t0[q].w += t1[q].w;
                }
                for (int sp = 2; sp < nslots; ++sp) {  // (small row blocks against many columns)
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int64_t r = min(rb + q, a_count - 1);
                        t1[q] = zero4;
                        if (k < D) t1[q] = __ldcg(reinterpret_cast<const float4*>(acc_out + ((int64_t)sp * a_count + r) * D + k));
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        t0[q].x += t1[q].x; t0[q].y += t1[q].y; t0[q].z += t1[q].z; t0[q].w += t1[q].w;
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int64_t r = rb + q;
                    float4 du;
                    du.x = (-tail.pos_coeff * up[q].x + t0[q].x) * tail.scale;
                    du.y = (-tail.pos_coeff * up[q].y + t0[q].y) * tail.scale;
                    du.z = (-tail.pos_coeff * up[q].z + t0[q].z) * tail.scale;
                    du.w = (-tail.pos_coeff * up[q].w + t0[q].w) * tail.scale;
                    float4 uu4 = zz[q];
                    uu4.x *= inv[q]; uu4.y *= inv[q]; uu4.z *= inv[q]; uu4.w *= inv[q];
                    const float dot = warp_sum(du.x * uu4.x + du.y * uu4.y + du.z * uu4.z + du.w * uu4.w);
                    const float gi = g * inv[q];
                    if (r < a_count && k < D) {
                        const float4 o = make_float4(gi * (du.x - dot * uu4.x), gi * (du.y - dot * uu4.y),
                                                     gi * (du.z - dot * uu4.z), gi * (du.w - dot * uu4.w));
                        if (tail.z_bf16) {
                            uint2 pk;
                            pk.x = pack_bf16(o.x, o.y);
                            pk.y = pack_bf16(o.z, o.w);
                            *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(tail.dz_out) + r * D + k) = pk;
                        } else {
                            *reinterpret_cast<float4*>(static_cast<float*>(tail.dz_out) + r * D + k) = o;
                        }
                    }
                }
            }
        };

        bool have = bwd_next_unit(t_next, t_end, col_tiles, total_tiles, ctas, u);
        if (have) park_q(u.rt);
        int cb = 0;  // column-constant blocks consumed so far
        while (have) {
            const int64_t a0 = a_begin + (int64_t)u.rt * kTile;
            const int64_t row = a0 + row_in_tile;
            const bool row_ok = row < a_end;
            const int jt0 = u.ct0, n_tiles = u.n;
            // column tiles that need the masked weight path, as 32-bit tile numbers: those that
            // overlap this row tile's own rows (the diagonal) and a ragged last tile
            const int64_t ub = unit_b_begin(u.rt), b_end = ub + b_count;
            const int64_t dd = a0 - ub;
            const int jt_diag_lo = (int)(dd >= 0 ? dd / kTile : -((-dd + kTile - 1) / kTile));
            const int jt_diag_hi = (int)(dd >= 0 ? (dd + kTile - 1) / kTile : -((-dd) / kTile));
            const int jt_ragged = (b_count % kTile) ? col_tiles - 1 : -1;
            NTX_TRACE_U(un, 0, clock64());
            NTX_TRACE_U(un, 7, (unsigned long long)n_tiles);

            // per-row constant: l_a in base-2 units, or 2^-l_a for the one-exponential form
            float row_c = 0.f;
            if (MODE != kCol) {
                if (row_ok) {
                    const float l2 = ld_f1_coherent(lse + row) * kLog2e;  // this rank's own rows
                    row_c = (MODE == kSym && !kTwoExp) ? ex2(-l2) : l2;
                } else {
                    row_c = (MODE == kSym && !kTwoExp) ? 0.f : INFINITY;  // rows past the range: W = 0
                }
            }
            // The tile loop.  A tile's barriers, TMEM columns and phase bits depend on g & 1 (and on
            // g >> 1, which flips every other tile): the body exists once per parity so that all of
            // them are immediates.  The rare masked tiles (diagonal, ragged edge) and the hand-over of
            // the column-constant blocks are kept out of the steady-state loop: every reconvergence
            // point and every dependent address computation in it was 30-50 clk of a 1550 clk tile
            // (profiles/r2_ncu_bwd_summary.csv and the source view behind it).
            int it = 0;
            uint32_t half = ((uint32_t)g >> 1) & 1u;  // (g >> 1) & 1, kept incrementally
            const float* cc = colc;
            auto tile = [&](auto par_c, auto special_c) {
                constexpr int PAR = decltype(par_c)::value;  // g & 1
                constexpr bool kSpecial = decltype(special_c)::value;
                constexpr int sb = B::kSBufs == 2 ? PAR : 0, pb = B::kPBufs == 2 ? PAR : 0;
                const uint32_t s_par = B::kSBufs == 2 ? half : (uint32_t)PAR;       // (g / kSBufs) & 1
                const uint32_t p_par = (B::kPBufs == 2 ? half : (uint32_t)PAR) ^ 1u;  // W block free again
                const uint32_t taddr = lane_addr + sb * 128 + quarter * 32;  // this warp's 32 S columns
                uint32_t v[32], packed[16];
                (void)packed;
                (void)p_par;
                NTX_TRACE(trole, it, 0);
                mbar_wait(&s_full[sb], s_par);
                tc_fence_after();
                NTX_TRACE(trole, it, 1);
                if (it == 0) NTX_TRACE_U(un, 1, clock64());
                tmem_ld32(taddr, v);
                // two W blocks: block pb was last read by the second MMA of tile g-2, issued a whole
                // tile ago: checked here so that the barrier's latency hides under the TMEM load
                if (SEP && B::kPBufs == 2) mbar_wait(&p_free[pb], p_par);
                tmem_ld_wait();
                if (SEP) {  // S is in registers: the buffer can take the S tile after the next
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&s_free[sb]);
                }
                NTX_TRACE(trole, it, 2);
                const int64_t j0 = ub + (int64_t)(jt0 + it) * kTile + quarter * 32;  // (masked path only)
                if (TF32)  // 32-bit weights, in place over the S values
                    weight_chunk<MODE, kTwoExp, kSpecial, false, 32>(v, v, cc, k_scale, row_c, j0, row, b_end);
                else
                    weight_chunk<MODE, kTwoExp, kSpecial>(v, packed, cc, k_scale, row_c, j0, row, b_end);
#ifdef NTXENT_DROP_TILE
                // Fault injection for the parity tests' own test (tools/drop_tile_proof.py): a build
                // with this define loses the weights of ONE tile; the full-size comparison must notice.
                if (u.rt == 100 && jt0 + it == 37) {
#pragma unroll
                    for (int e = 0; e < 16; ++e) packed[e] = 0u;
                }
#endif
                NTX_TRACE(trole, it, 3);
                if (SEP) {
                    // one W block: the second MMA of the previous tile, issued while this tile's
                    // weights were being computed, must have read it
                    if (B::kPBufs == 1) mbar_wait(&p_free[0], p_par);
                    tc_fence_after();
                    NTX_TRACE(trole, it, 4);
                    tmem_st16(lane_addr + B::kPCol + pb * 64 + quarter * 16, packed);
                } else if (TF32) {
                    tmem_st32(taddr, v);  // 32 tf32 weights over the 32 S columns just consumed
                } else {
                    tmem_st16(taddr, packed);  // over the S columns just consumed
                }
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&p_full[pb]);
                NTX_TRACE(trole, it, 5);
                ++it;
                ++g;
                cc += kTile;
                if (PAR) half ^= 1u;
            };
            using Even = std::integral_constant<int, 0>;
            using Odd = std::integral_constant<int, 1>;
            // masked tiles, as positions in this unit's walk
            const int sp_lo = jt_diag_lo - jt0, sp_hi = jt_diag_hi - jt0, sp_rg = jt_ragged - jt0;
            while (it < n_tiles) {
                // one block of column constants (staged by the constants warp)
                const int blk_end = min(n_tiles, it + B::kCcTiles);
                if (MODE != kRow) mbar_wait(&cc_full[cb & 1], (cb >> 1) & 1);
                cc = colc + (cb & 1) * B::kCcTiles * kTile + quarter * 32;
                while (it < blk_end) {
                    int ns = blk_end;  // the next masked tile of the block
                    if (it <= sp_hi) ns = min(ns, max(it, sp_lo));
                    if (sp_rg >= it) ns = min(ns, sp_rg);
                    int n = ns - it;
                    if (n > 0 && (g & 1)) {
                        tile(Odd{}, std::false_type{});
                        --n;
                    }
                    for (; n >= 2; n -= 2) {
                        tile(Even{}, std::false_type{});
                        tile(Odd{}, std::false_type{});
                    }
                    if (n) tile(Even{}, std::false_type{});
                    if (it < blk_end) {  // it == ns: a masked tile
                        if (g & 1)
                            tile(Odd{}, std::true_type{});
                        else
                            tile(Even{}, std::true_type{});
                    }
                }
                if (MODE != kRow) {
                    if (lane == 0) mbar_arrive(&cc_empty[cb & 1]);  // (after the last tile's __syncwarp)
                    ++cb;
                }
            }
            // The S issuer can start on the next unit while this one's last W U runs and its
            // accumulator is read out: hand it the next row tile first.
            BwdUnit nu;
            const bool have_next = bwd_next_unit(t_next, t_end, col_tiles, total_tiles, ctas, nu);
            NTX_TRACE_U(un, 2, clock64());
            if (have_next) park_q(nu.rt);
            NTX_TRACE_U(un, 3, clock64());

            // unit epilogue: raw accumulator rows -> acc_out[slot][row][D], 32-column groups
            // round-robin; the accumulator is handed back as soon as the last group is in registers
            mbar_wait(acc_full, un & 1);
            tc_fence_after();
            NTX_TRACE_U(un, 4, clock64());
            const int64_t r = (int64_t)u.rt * kTile + row_in_tile;
            float* dst = acc_out + ((int64_t)u.slot * a_count + r) * D;
            auto release_acc = [&] {
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_free);
            };
            if (quarter >= D / 32) release_acc();
            for (int c = quarter; c < D / 32; c += 4) {
                uint32_t v[32];
                tmem_ld32(lane_addr + B::kAccCol + c * 32, v);
                tmem_ld_wait();
                if (c + 4 >= D / 32) release_acc();
                if (r < a_count) {
#pragma unroll
                    for (int e = 0; e < 32; e += 4)
                        *reinterpret_cast<float4*>(dst + c * 32 + e) =
                            make_float4(__uint_as_float(v[e]), __uint_as_float(v[e + 1]),
                                        __uint_as_float(v[e + 2]), __uint_as_float(v[e + 3]));
                }
            }
            NTX_TRACE_U(un, 5, clock64());
            if (!TF32 && D <= 128 && tail.tickets) tail_unit(u);
            u = nu;
            have = have_next;
            ++un;
        }
    }
    tc_fence_before();
    __syncthreads();
    NTX_TRACE_U(63, 2, clock64());
    NTX_TRACE_U(63, 7, globaltimer_ns());
    if (warp == B::kMmaWarp) tmem_dealloc<512>(tmem_base);
}

// --------------------------------------------------------------------- host ---
PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPointByVersion("cuTensorMapEncodeTiled", &p, 12000, cudaEnableDefault, &q) ==
                cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    });
    return fn;
}

// [M, D] row-major (bf16 or fp32), box = [128 rows x 128 bytes], 128-byte swizzle, zero fill past
// the edge.
cudaError_t make_tmap_2d(const void* base, bool tf32, int64_t rows, int64_t cols, int box_rows,
                         CUtensorMap* out) {
    auto fn = encode_fn();
    if (!fn) return cudaErrorNotSupported;
    const void* u_all = base;
    const int64_t M = rows, D = cols;
    const int eb = tf32 ? 4 : 2;
    cuuint64_t dims[2] = {(cuuint64_t)D, (cuuint64_t)M};
    cuuint64_t strides[1] = {(cuuint64_t)D * eb};
    cuuint32_t box[2] = {(cuuint32_t)(128 / eb), (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    auto encode = [&] {
        return fn(out, tf32 ? CU_TENSOR_MAP_DATA_TYPE_TFLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(u_all), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    };
    CUresult r = encode();
    if (r == CUDA_ERROR_INVALID_CONTEXT) {
        // a host thread that has not touched the runtime yet (an autograd worker thread, say) has
        // no driver context bound: let the runtime bind the current device's primary context
        (void)cudaFree(nullptr);
        r = encode();
    }
    if (r != CUDA_SUCCESS && getenv("NTXENT_DEBUG"))
        fprintf(stderr, "[ntxent_b200] cuTensorMapEncodeTiled(ptr=%p, M=%lld, D=%lld) -> CUresult %d\n", u_all,
                (long long)M, (long long)D, (int)r);
    return r == CUDA_SUCCESS ? cudaSuccess : cudaErrorInvalidValue;
}

cudaError_t make_tmap(const void* u_all, bool tf32, int64_t M, int64_t D, CUtensorMap* out) {
    return cached_tmap_2d(u_all, tf32, M, D, kTile, out);
}

template <int D, bool TF32, bool kOnline>
cudaError_t launch_fwd_t(const CUtensorMap& tmap, int64_t M, int64_t row_begin, int64_t row_count,
                         float k_scale, const SplitPlan& plan, int col_tiles, int cross_half, float* part_m,
                         float* part_s, const PeerWait& u_ready, const void* u_raw, const FwdTail& tail,
                         cudaStream_t stream) {
    using C = Cfg<D, TF32>;
    // request enough shared memory that no more CTAs than the TMEM budget allows co-reside
    const int smem = C::kFwdCtasPerSm == 2 ? max(C::kFwdSmem, 80 * 1024) : max(C::kFwdSmem, 120 * 1024);
    auto kern = tc_fwd_kernel<D, TF32, kOnline>;
    static unsigned long long attr_done = 0;
    cudaError_t e = set_smem_once(kern, smem, attr_done);
    if (e != cudaSuccess) return e;
    dim3 grid((unsigned)plan.row_tiles, (unsigned)plan.nsplit);
    return launch_kernel(kern, grid, dim3(kThreads), smem, stream, tmap, M, row_begin, row_count, k_scale,
                         (int)(plan.cols_per_split / kTile), col_tiles, cross_half, part_m, part_s, u_ready, u_raw,
                         tail);
}

// U [M, D] fp32 -> UT [D, Mpad] (Mpad = M rounded up to 4, pad columns zero): the K-major B operand
// of the tf32 second MMA.  32x32 shared-memory transpose; 2 x 4 MB of traffic at 8192 x 128.
__global__ void __launch_bounds__(256) transpose_f32_kernel(const float* __restrict__ U, int64_t M, int D,
                                                            int64_t Mpad, float* __restrict__ UT) {
    __shared__ float tile[32][33];
    pdl_trigger();
    pdl_wait();
    const int64_t r0 = (int64_t)blockIdx.x * 32;
    const int d0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    for (int i = ty; i < 32; i += 8) tile[i][tx] = (r0 + i < M) ? U[(r0 + i) * D + d0 + tx] : 0.f;
    __syncthreads();
    for (int i = ty; i < 32; i += 8)
        if (r0 + tx < Mpad) UT[(int64_t)(d0 + i) * Mpad + r0 + tx] = tile[tx][i];
}

template <int D, bool TF32, int MODE, bool kTwoExp>
cudaError_t launch_bwd_t(const CUtensorMap& tmap, const CUtensorMap& tmap_t, const void* u_all,
                         const float* lse, int64_t M,
                         int64_t a_begin, int64_t a_count, int64_t b_begin, int64_t b_count,
                         float k_scale, const SplitPlan& plan, int col_tiles, int cross_half, float* acc,
                         const PeerWait& lse_ready, const BwdTail& tail, cudaStream_t stream) {
    using B = BwdCfg<D, TF32>;
    const int smem = max(B::kSmem, 120 * 1024);
    auto kern = tc_bwd_kernel<D, TF32, MODE, kTwoExp>;
    static unsigned long long attr_done = 0;
    cudaError_t e = set_smem_once(kern, smem, attr_done);
    if (e != cudaSuccess) return e;
    const int64_t total_tiles = (int64_t)plan.row_tiles * col_tiles;
    if (plan.sk_ctas < 1 || plan.sk_ctas > total_tiles) return cudaErrorInvalidValue;
    return launch_kernel(kern, dim3((unsigned)plan.sk_ctas), dim3(B::kThreads), smem, stream, tmap, tmap_t,
                         u_all, lse, M, a_begin, a_count, b_begin, b_count, k_scale, col_tiles, total_tiles,
                         cross_half, acc, g_trace, lse_ready, tail);
}

template <int D, bool TF32>
cudaError_t launch_bwd_d(const CUtensorMap& tmap, const CUtensorMap& tmap_t, const void* u_all,
                         const float* lse, int64_t M,
                         int64_t a_begin, int64_t a_count, int64_t b_begin, int64_t b_count,
                         float k_scale, int mode, bool two_exp, const SplitPlan& plan, int col_tiles,
                         int cross_half, float* acc, const PeerWait& lse_ready, const BwdTail& tail,
                         cudaStream_t stream) {
#define NTX_BWD(MODE, TWO)                                                                               \
    launch_bwd_t<D, TF32, MODE, TWO>(tmap, tmap_t, u_all, lse, M, a_begin, a_count, b_begin, b_count, k_scale, \
                                     plan, col_tiles, cross_half, acc, lse_ready, tail, stream)
    if (mode == kRow) return NTX_BWD(kRow, false);
    if (mode == kCol) return NTX_BWD(kCol, false);
    return two_exp ? NTX_BWD(kSym, true) : NTX_BWD(kSym, false);
#undef NTX_BWD
}

}  // namespace

// Encoding a tensor map is ~1-2 us of host time per call and the same few (pointer, shape) pairs
// come back every step: keep the last ones.  (A map only encodes address, shape and layout, so a
// hit stays valid when the buffer was freed and handed out again at the same address.)
struct TmapKey {
    const void* base; int64_t rows, cols; int box_rows; bool tf32;
    bool operator==(const TmapKey& o) const {
        return base == o.base && rows == o.rows && cols == o.cols && box_rows == o.box_rows && tf32 == o.tf32;
    }
};
cudaError_t tc::cached_tmap_2d(const void* base, bool tf32, int64_t rows, int64_t cols, int box_rows,
                           CUtensorMap* out) {
    constexpr int kSlots = 32;
    static std::mutex mu;
    static TmapKey keys[kSlots];
    static CUtensorMap maps[kSlots];
    static int used = 0, next = 0;
    const TmapKey key{base, rows, cols, box_rows, tf32};
    {
        std::lock_guard<std::mutex> lock(mu);
        for (int i = 0; i < used; ++i)
            if (keys[i] == key) { *out = maps[i]; return cudaSuccess; }
    }
    cudaError_t e = make_tmap_2d(base, tf32, rows, cols, box_rows, out);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    const int slot = used < kSlots ? used++ : (next++ % kSlots);
    keys[slot] = key;
    maps[slot] = *out;
    return cudaSuccess;
}


cudaError_t tc::set_smem_once_impl(const void* kern, int smem, unsigned long long& done_mask) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (dev < 64 && ((done_mask >> dev) & 1ull)) return cudaSuccess;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e == cudaSuccess && dev < 64) done_mask |= 1ull << dev;
    return e;
}

bool tc_set_trace_buffer(unsigned long long* dev_ptr) {
    if (dev_ptr && !NTXENT_TRACE) return false;  // this build has no time stamps compiled in
    g_trace = dev_ptr;
    return true;
}

int64_t tc_ut_cols(int64_t M) { return (M + 3) / 4 * 4; }

size_t tc_bwd_scratch_bytes(int64_t M, int64_t D, bool op_bf16) {
    return op_bf16 ? 0 : (size_t)tc_ut_cols(M) * D * sizeof(float);
}

bool tc_bwd_tail_supported(int64_t D, bool op_bf16) { return op_bf16 && (D == 64 || D == 128); }

bool tc_supported(int64_t M, int64_t D, bool op_bf16) {
    if (op_bf16) return D == 64 || D == 128 || D == 192 || D == 256;
    // fp32 operands run as tf32 on the tensor cores once the problem is large enough for speed to
    // matter; below that (and for D the tf32 tiles do not cover) the exact fp32-FMA kernels run
    return (D == 64 || D == 128) && M >= kTf32MinRows;
}

SplitPlan tc_plan_fwd(int64_t M, int64_t D, int64_t row_count, bool op_bf16, int num_sms) {
    const int tile_bytes = (int)(D * (op_bf16 ? 2 : 4) * kTile);
    const int stages = tile_bytes <= 16 * 1024 ? 3 : 2;
    const int per_sm = ((1 + stages) * tile_bytes + 2048 <= 112 * 1024) ? 2 : 1;
    return plan_splits(row_count, M, kTile, kTile, num_sms * per_sm, 32);
}

// One persistent CTA per SM, each with an equal share of the tiles (see BwdUnit).  nsplit = the
// largest number of CTAs whose ranges touch one row tile = partial slots the epilogue may have to add.
SplitPlan tc_plan_bwd(int64_t M, int64_t D, int64_t a_count, int64_t b_count, int num_sms) {
    (void)M; (void)D;
    SplitPlan p{};
    p.row_tiles = (int)((a_count + kTile - 1) / kTile);
    const int64_t col_tiles = std::max<int64_t>(1, (b_count + kTile - 1) / kTile);
    const int64_t total = (int64_t)p.row_tiles * col_tiles;
    p.sk_ctas = (int)std::min<int64_t>(std::max(num_sms, 1), total);
    p.sk_col_tiles = (int)col_tiles;
    p.cols_per_split = col_tiles * kTile;
    p.nsplit = 1;
    for (int rt = 0; rt < p.row_tiles; ++rt) {
        const int64_t first = (int64_t)rt * col_tiles;
        const int n = (int)(bwd_cta_of_tile(first + col_tiles - 1, total, p.sk_ctas) -
                            bwd_cta_of_tile(first, total, p.sk_ctas)) + 1;
        p.nsplit = std::max(p.nsplit, n);
    }
    return p;
}

cudaError_t tc_launch_fwd(const void* u_all, bool op_bf16, int64_t M, int64_t D, int64_t row_begin,
                          int64_t row_count, float tau, unsigned flags, const SplitPlan& plan,
                          float* part_m, float* part_s, const PeerWait& u_ready,
                          cudaStream_t stream, bool cross_view, const FwdTail* tail_in) {
    FwdTail tail{};
    if (tail_in) tail = *tail_in;
    CUtensorMap tmap;
    cudaError_t e = make_tmap(u_all, !op_bf16, M, D, &tmap);
    if (e != cudaSuccess) return e;
    const float k_scale = kLog2e / tau;
    const bool online = (flags & 4u) || k_scale > kFixedMaxLimit;
    // cross-view form: every row tile walks the T/2 tiles of the other view (M/2 a multiple of 128)
    if (cross_view && ((M / 2) % kTile != 0 || u_ready.flags)) return cudaErrorInvalidValue;
    const int cross_half = cross_view ? (int)(M / 2 / kTile) : 0;
    const int col_tiles = cross_view ? cross_half : (int)((M + kTile - 1) / kTile);
#define NTX_FWD(DD, TF)                                                                                    \
    (online ? launch_fwd_t<DD, TF, true>(tmap, M, row_begin, row_count, k_scale, plan, col_tiles, cross_half,  \
                                         part_m, part_s, u_ready, u_all, tail, stream)                         \
            : launch_fwd_t<DD, TF, false>(tmap, M, row_begin, row_count, k_scale, plan, col_tiles, cross_half, \
                                          part_m, part_s, u_ready, u_all, tail, stream))
    if (!op_bf16) {
        switch (D) {
            case 64: return NTX_FWD(64, true);
            case 128: return NTX_FWD(128, true);
            default: return cudaErrorNotSupported;
        }
    }
    switch (D) {
        case 64: return NTX_FWD(64, false);
        case 128: return NTX_FWD(128, false);
        case 192: return NTX_FWD(192, false);
        case 256: return NTX_FWD(256, false);
        default: return cudaErrorNotSupported;
    }
#undef NTX_FWD
}

cudaError_t tc_launch_bwd(const void* u_all, bool op_bf16, const float* lse_all, int64_t M, int64_t D,
                          int64_t a_begin, int64_t a_count, int64_t b_begin, int64_t b_count,
                          float tau, int mode, unsigned flags, const SplitPlan& plan, float* acc,
                          float* ut_scratch, const PeerWait& lse_ready, cudaStream_t stream, bool cross_view,
                          const BwdTail* tail_in) {
    BwdTail tail{};
    if (tail_in) {
        if (!tc_bwd_tail_supported(D, op_bf16) || mode != kSym || !tail_in->tickets) return cudaErrorInvalidValue;
        tail = *tail_in;
    }
    // cross-view form: the caller passes b_begin = 0, b_count = M/2 (a multiple of 128) and mode kSym
    if (cross_view && ((M / 2) % kTile != 0 || b_begin != 0 || b_count != M / 2 || mode != kSym || lse_ready.flags))
        return cudaErrorInvalidValue;
    const int cross_half = cross_view ? (int)(M / 2 / kTile) : 0;
    CUtensorMap tmap, tmap_t;
    cudaError_t e = make_tmap(u_all, !op_bf16, M, D, &tmap);
    if (e != cudaSuccess) return e;
    tmap_t = tmap;  // unused by the bf16 kernels
    if (!op_bf16) {
        if (!ut_scratch) return cudaErrorInvalidValue;
        const int64_t Mpad = tc_ut_cols(M);
        dim3 tgrid((unsigned)((Mpad + 31) / 32), (unsigned)(D / 32));
        e = launch_kernel(transpose_f32_kernel, tgrid, dim3(256), 0, stream, (const float*)u_all, M, (int)D,
                          Mpad, ut_scratch);
        if (e != cudaSuccess) return e;
        e = cached_tmap_2d(ut_scratch, true, /*rows=*/D, /*cols=*/Mpad, /*box_rows=*/D, &tmap_t);
        if (e != cudaSuccess) return e;
    }
    const float k_scale = kLog2e / tau;
    const bool two_exp = (flags & 4u) || k_scale > kFixedMaxLimit;
    const int col_tiles = (int)((b_count + kTile - 1) / kTile);
#define NTX_BWD_D(DD, TF)                                                                                   \
    launch_bwd_d<DD, TF>(tmap, tmap_t, u_all, lse_all, M, a_begin, a_count, b_begin, b_count, k_scale, mode, \
                         two_exp, plan, col_tiles, cross_half, acc, lse_ready, tail, stream)
    if (!op_bf16) {
        switch (D) {
            case 64: return NTX_BWD_D(64, true);
            case 128: return NTX_BWD_D(128, true);
            default: return cudaErrorNotSupported;
        }
    }
    switch (D) {
        case 64: return NTX_BWD_D(64, false);
        case 128: return NTX_BWD_D(128, false);
        case 192: return NTX_BWD_D(192, false);
        case 256: return NTX_BWD_D(256, false);
        default: return cudaErrorNotSupported;
    }
#undef NTX_BWD_D
}

}  // namespace ntxent_b200
