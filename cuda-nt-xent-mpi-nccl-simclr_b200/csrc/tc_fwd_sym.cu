// tc_fwd_sym.cu -- the forward tile pass over the UPPER TRIANGLE of the logit matrix.
//
// With the fixed-max form (every row uses the same offset, so nothing is rescaled)
// E_ab = 2^(k u_a.u_b) is symmetric: a tile (i, j), i < j, yields the row sums of tile i's rows
// AND, summed down its columns, the row sums of tile j's rows.  The kernel therefore visits
// T (T + 1) / 2 tiles instead of T^2: half the tensor-core work and half the exponentials of
// tc_fwd_kernel (tc_kernels.cu), which stays for everything this one does not cover (running-max
// logsumexp for cold temperatures, row blocks of a partitioned batch, small problems).
//
// Work items: row tile i owns the tiles (i, i .. T-1); they are cut into chunks of at most L
// tiles, one CTA per chunk, long rows first.  Per chunk a CTA keeps the row sums of its 128 rows
// in registers (as tc_fwd_kernel does) and emits, per tile, the column sums of each 32-row
// quadrant:  cpart[(i, j)][quadrant][column]  -- a fixed slot per contribution, so the finalize
// kernel adds them in a fixed order (deterministic, no atomics).
//
// The S tile is read out of TMEM with the 16x256b shape instead of 32x32b: a thread then holds
// 4 rows x 8 columns of every 32 x 32 block rather than one row x 32 columns, so the column sums
// need three shuffle levels (over the 8 lanes that share the columns) instead of five, and the
// row sums stay per-thread until the end of the chunk.
//
// Replaces, like tc_fwd_kernel, the reference's cublasSgemm + row_max/softmax kernels
// (src/ntxent_kernel.cu:8-103,162-199); the reference forms the full [2B,2B] matrix.
#include <algorithm>

#include "tc_kernels.cuh"
#include "tc_tile.cuh"

namespace ntxent_b200 {
using namespace ptx;
using namespace tc;

namespace {

constexpr int kSymThreads = 320;  // warps 0..7 softmax, warp 8 TMA, warp 9 MMA
constexpr int kSymTmaWarp = 8, kSymMmaWarp = 9;
constexpr int kSymSoftmaxThreads = 256;

// 16 TMEM lanes x 16 columns -> 8 registers: thread t holds rows t/4 (regs 4n, 4n+1) and t/4 + 8
// (regs 4n+2, 4n+3) of the 16, columns 8n + 2 (t%4) + {0, 1}, n = 0..1.
__device__ __forceinline__ void tmem_ld_16x256b_x2(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x2.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
        : "r"(taddr)
        : "memory");
}
// Work items.  Rows i in [T - m L, T - (m-1) L) have between (m-1) L + 1 and m L tiles, i.e. m
// chunks; items are numbered by descending m (ascending i), chunks of a row consecutively.
struct SymItem {
    int i, j0, n, slot;
};
__host__ __device__ inline int sym_item_count(int T, int L) {
    int total = 0;
    for (int m = (T + L - 1) / L; m >= 1; --m) {
        const int lo = max(T - m * L, 0), hi = T - (m - 1) * L;
        total += (hi - lo) * m;
    }
    return total;
}
__device__ __forceinline__ SymItem sym_item(int w, int T, int L) {
    SymItem it{0, 0, 0, 0};
    for (int m = (T + L - 1) / L; m >= 1; --m) {
        const int lo = max(T - m * L, 0), hi = T - (m - 1) * L;
        const int ctas = (hi - lo) * m;
        if (w < ctas) {
            it.i = lo + w / m;
            it.slot = w % m;
            it.j0 = it.i + it.slot * L;
            it.n = min(L, T - it.j0);
            return it;
        }
        w -= ctas;
    }
    return it;
}

template <int D, bool TF32>
__global__ void __launch_bounds__(kSymThreads, Cfg<D, TF32>::kFwdCtasPerSm)
tc_fwd_sym_kernel(const __grid_constant__ CUtensorMap tmap, int64_t M, float k_scale, int T, int L,
                  float* __restrict__ part_s, float* __restrict__ cpart, float* __restrict__ pos_t_out) {
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
    const SymItem item = sym_item((int)blockIdx.x, T, L);
    const int64_t a0 = (int64_t)item.i * kTile;
    const int n_tiles = item.n;

    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmap);
        mbar_init(q_full, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&s_full[b], 1); mbar_init(&s_empty[b], 8); }
        mbar_init(q_ready, 8);
        fence_mbar_init();
    }
    if (warp == kSymMmaWarp) tmem_alloc<256>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();  // set-up done; from here on the kernel reads what the normalise kernel wrote

    if (warp == kSymTmaWarp) {
        if (lane == 0) {
            mbar_arrive_expect_tx(q_full, C::kTileBytes);
            for (int a = 0; a < C::kAtoms; ++a)
                tma_load_2d(q_tile + a * kAtomBytes, &tmap, q_full, a * C::kAtomCols, (int32_t)a0);
            for (int it = 0; it < n_tiles; ++it) {
                const int st = it % NS;
                mbar_wait(&k_empty[st], ((it / NS) & 1) ^ 1);
                mbar_arrive_expect_tx(&k_full[st], C::kTileBytes);
                for (int a = 0; a < C::kAtoms; ++a)
                    tma_load_2d(k_tiles + st * C::kTileBytes + a * kAtomBytes, &tmap, &k_full[st],
                                a * C::kAtomCols, (item.j0 + it) * kTile);
            }
        }
        __syncwarp();
    } else if (warp == kSymMmaWarp) {
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
        // eight softmax warps: warp w owns TMEM lanes 32 (w%4) .. +31 and the 64 columns
        // [64 h, 64 h + 64) of every S tile, h = w/4, as two 32 x 32 blocks
        const int quad = warp & 3;
        const int half = warp >> 2;
        const uint32_t quad_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + half * 64;
        // this thread's four rows (tile-relative): rbase + {0, 8, 16, 24}; its columns within a
        // 32-column block: 8 n + cbase + {0, 1}
        const int rbase = quad * 32 + (lane >> 2);
        const int cbase = 2 * (lane & 3);
        // row sums, as packed pairs (even column, odd column) of the thread's four rows
        f32x2 rs2[4] = {0ull, 0ull, 0ull, 0ull};
        // scale the row tile by k = log2(e)/tau where TMA put it (elementwise, so the swizzle is
        // irrelevant), then publish it to the tensor core's (async-proxy) view of shared memory
        {
            mbar_wait(q_full, 0);
            uint4* q4 = reinterpret_cast<uint4*>(q_tile);
            for (int i = threadIdx.x; i < C::kTileBytes / 16; i += kSymSoftmaxThreads) {
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
        // 64-column half tiles that take the masked path, as 32-bit numbers: the diagonal tile, the
        // ragged end of the range, and the half tiles that hold the positives of this row tile
        // (columns (a0 + N + r) mod M: one run of 128 columns, or two when it wraps around M)
        const int hj_ragged = (int)(M / 64);
        const int p0 = (int)pair_row(a0, M);
        const int hj_pos_lo = p0 / 64, hj_pos_hi = (min(p0 + kTile, (int)M) - 1) / 64;
        const int hj_wrap_hi = p0 + kTile > (int)M ? (p0 + kTile - (int)M - 1) / 64 : -1;
        // column partials of tile (i, j): cpart[(j (j-1)/2 + i) * 512 + quad * 128 + column]
        // (all row / column numbers fit 32 bits: M <= 2^24)
        const int M32 = (int)M, N32 = (int)(M / 2), row0 = (int)a0 + rbase;
        int pos4[4];  // the positives of this thread's four rows
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const int p = row0 + 8 * g + N32;
            pos4[g] = p >= M32 ? p - M32 : p;
        }
        // The column sums are transposed while they are reduced: at each of three levels a lane
        // keeps one half of its values and hands the other half to its partner (lane ^ 16, 8, 4).
        // Which half is a bit of the lane number; as all-ones / all-zeros words the choice is one
        // LOP3 per value (a & m | b & ~m) with three registers held across the loop -- as predicates
        // the compiler rebuilt them from the lane number in every block (seen in the ncu source view).
        const uint32_t m4 = (lane & 16) ? 0xffffffffu : 0u, m3 = (lane & 8) ? 0xffffffffu : 0u,
                       m2 = (lane & 4) ? 0xffffffffu : 0u;
        auto pick = [](uint32_t m, float a, float b) {  // m ? a : b
            return __uint_as_float((__float_as_uint(a) & m) | (__float_as_uint(b) & ~m));
        };
        // cs[2 n + b] = this thread's sum (4 rows) of column 8 n + cbase + b of a 16-column block ->
        // the sum over lane bits 4 and 3 of column 8 b4 + cbase + b3 (bit 2 still to be added)
        auto reduce43 = [&](const float (&cs)[4]) -> float {
            const float k0 = pick(m4, cs[2], cs[0]) + __shfl_xor_sync(0xffffffffu, pick(m4, cs[0], cs[2]), 16);
            const float k1 = pick(m4, cs[3], cs[1]) + __shfl_xor_sync(0xffffffffu, pick(m4, cs[1], cs[3]), 16);
            return pick(m3, k1, k0) + __shfl_xor_sync(0xffffffffu, pick(m3, k0, k1), 8);
        };
        // Two neighbouring blocks finish together: the last level (lane bit 2) hands block A's column
        // to the lanes with b2 = 0 and block B's to those with b2 = 1 -- one shuffle, and the warp
        // stores 32 finished column sums (one 128-byte line) of its quadrant.
        auto finish2 = [&](float pa, float pb) -> float {
            return pick(m2, pb, pa) + __shfl_xor_sync(0xffffffffu, pick(m2, pa, pb), 4);
        };
        // this lane's slot in the 512 column partials of a tile pair: quadrant, column half, and
        // column 16 b2 + 8 b4 + cbase + b3 of the 32 that a pair of blocks covers
        const int col_of_lane = ((lane & 4) ? 16 : 0) + ((lane & 16) ? 8 : 0) + cbase + ((lane & 8) ? 1 : 0);
        // cpart[(j (j-1)/2 + i) * 512 + ...] for the tile j of this iteration: kept as a pointer and
        // advanced by j * 512 from one tile to the next (pair(j+1) - pair(j) = j)
        float* cp = cpart + ((int64_t)item.j0 * (item.j0 - 1) / 2 + item.i) * 512 + quad * 128 + half * 64 +
                    col_of_lane;

        // One 32-row x 16-column block on the plain path: exponentials, row sums (packed, into rs2)
        // and column sums of the thread's four rows.
        auto block_fast = [&](const uint32_t (&v)[16]) -> float {
            // registers r, r+1 are one row's two neighbouring columns: one packed add each into
            // the row pair and the column pair
            f32x2 cs2[2] = {0ull, 0ull};
#pragma unroll
            for (int r = 0; r < 16; r += 2) {
                const f32x2 e = ex2_pair<kFwdPolyPairs>(v[r], v[r + 1], r >> 1);
                const int g = 2 * (r >> 3) + ((r >> 1) & 1);
                rs2[g] = add2(rs2[g], e);
                cs2[(r >> 2) & 1] = add2(cs2[(r >> 2) & 1], e);
            }
            float cs[4];
            unpack2(cs2[0], cs[0], cs[1]);
            unpack2(cs2[1], cs[2], cs[3]);
            return reduce43(cs);
        };
        // The same on the masked path (diagonal tile, ragged edge, the tiles with the positives).
        auto block_slow = [&](const uint32_t (&v)[16], int jt, int blk) -> float {
            const bool diag = jt == item.i;
            const int c0 = jt * kTile + half * 64 + blk * 16 + cbase;
            float rs[4] = {0.f, 0.f, 0.f, 0.f}, cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int g = 2 * (r >> 3) + ((r >> 1) & 1);
                const int row = row0 + 8 * g;
                const int col = c0 + 8 * ((r >> 2) & 1) + (r & 1);
                const float t = __uint_as_float(v[r]);
                const bool ok = col < M32 && col != row && row < M32;
                const float e = ok ? ex2(t) : 0.f;
                rs[g] += e;
                cs[2 * ((r >> 2) & 1) + (r & 1)] += e;
                // the positive logit, from the same MMA as the denominator; off the diagonal tile
                // the element (row, col) is also the positive logit of row `col`
                if (ok && col == pos4[g]) {
                    pos_t_out[row] = t;
                    if (!diag) pos_t_out[col] = t;
                }
            }
#pragma unroll
            for (int g = 0; g < 4; ++g) rs2[g] = add2(rs2[g], pack2(rs[g], 0.f));
            return reduce43(cs);
        };
        auto ld_blk = [&](uint32_t col, uint32_t (&v)[16]) {
            tmem_ld_16x256b_x2(quad_addr + col, &v[0]);
            tmem_ld_16x256b_x2(quad_addr + (16u << 16) + col, &v[8]);
        };
        // the S buffer has been read out by this warp: hand it back and fetch the first block of the
        // next tile
        auto release_and_prefetch = [&](int it, uint32_t (&v)[16]) {
            const int buf = it & 1;
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[buf]);
            if (it + 1 < n_tiles) {
                mbar_wait(&s_full[buf ^ 1], ((it + 1) >> 1) & 1);
                tc_fence_after();
                ld_blk((buf ^ 1) * kTile, v);
            }
        };

        // four 16-column blocks per tile, double-buffered in registers: the load of the next block
        // (or of the next tile's first block) is in flight while the current one is exponentiated.
        // ONE branch per tile separates the plain path from the masked one.
        uint32_t va[16], vb[16];
        mbar_wait(&s_full[0], 0);
        tc_fence_after();
        ld_blk(0, va);
        for (int it = 0; it < n_tiles; ++it) {
            const int jt = item.j0 + it;
            const int hj = 2 * jt + half;
            const bool special = jt == item.i || hj >= hj_ragged || (hj >= hj_pos_lo && hj <= hj_pos_hi) ||
                                 hj <= hj_wrap_hi;
            const uint32_t tcol = (it & 1) * kTile;
            if (!special) {
                tmem_ld_wait();
                ld_blk(tcol + 16, vb);
                float pa = block_fast(va);
                tmem_ld_wait();
                ld_blk(tcol + 32, va);
                float pb = block_fast(vb);
                cp[0] = finish2(pa, pb);
                tmem_ld_wait();
                ld_blk(tcol + 48, vb);
                pa = block_fast(va);
                tmem_ld_wait();
                release_and_prefetch(it, va);
                pb = block_fast(vb);
                cp[32] = finish2(pa, pb);
            } else {
                // (rare: no double buffering, the blocks one after the other)
                float pa = 0.f;
                tmem_ld_wait();
#pragma unroll 1
                for (int blk = 0; blk < 4; ++blk) {
                    if (blk > 0) {
                        ld_blk(tcol + 16 * blk, va);
                        tmem_ld_wait();
                    }
                    const float pv = block_slow(va, jt, blk);
                    if (blk & 1) {
                        const float sum = finish2(pa, pv);
                        if (jt != item.i) cp[(blk >> 1) * 32] = sum;  // (a diagonal tile's column sums are its row sums)
                    } else {
                        pa = pv;
                    }
                }
                release_and_prefetch(it, va);
            }
            cp += (int64_t)jt * 512;
        }
        // row sums: over the four lanes that share the rows, then the two column halves through
        // shared memory (the row tile is no longer needed: every MMA that read it has completed)
        float rs[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            float lo, hi;
            unpack2(rs2[g], lo, hi);
            rs[g] = lo + hi;
            rs[g] += __shfl_xor_sync(0xffffffffu, rs[g], 1);
            rs[g] += __shfl_xor_sync(0xffffffffu, rs[g], 2);
        }
        float* halves = reinterpret_cast<float*>(q_tile);
        named_bar_sync(1, kSymSoftmaxThreads);
        if (half == 1 && (lane & 3) == 0) {
#pragma unroll
            for (int g = 0; g < 4; ++g) halves[rbase + 8 * g] = rs[g];
        }
        named_bar_sync(1, kSymSoftmaxThreads);
        if (half == 0 && (lane & 3) == 0) {
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const int64_t row = a0 + rbase + 8 * g;
                if (row < M) part_s[(int64_t)item.slot * M + row] = rs[g] + halves[rbase + 8 * g];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kSymMmaWarp) tmem_dealloc<256>(tmem_base);
}

// ---------------------------------------------------------------- finalize ---
// One CTA per row tile j, four threads per row: thread (g, c) adds the column partials of the
// tiles (i, j), i = g, g + 4, ... (eight loads in flight), the four are combined through shared
// memory in a fixed order, and thread (0, c) adds the row-sum partials (one per chunk of the row
// tile).  Fixed order throughout: run-to-run deterministic.
constexpr int kSymFinThreads = 512;
__global__ void __launch_bounds__(kSymFinThreads)
fwd_sym_finalize_kernel(int64_t M, int T, int L, const float* __restrict__ part_s,
                        const float* __restrict__ cpart, const float* __restrict__ pos_t,
                        float* __restrict__ lse_out, float* __restrict__ block_partials,
                        unsigned* __restrict__ counter, float* __restrict__ loss_out, float loss_scale) {
    __shared__ float part[3][kTile];
    __shared__ float warp_contrib[kTile / 32];
    __shared__ unsigned ticket_s;
    __shared__ double red[kSymFinThreads];
    pdl_trigger();
    pdl_wait();
    // row tile == column tile of these rows; the tiles with the most partials (large j) go first,
    // and sixteen loads are in flight per thread: the kernel was bound by the latency chain of its
    // longest block (16.3 us at C4 for 67 MB)
    const int j = (int)gridDim.x - 1 - (int)blockIdx.x;
    const int c = threadIdx.x & (kTile - 1), g = threadIdx.x >> 7;
    const int64_t r = (int64_t)j * kTile + c;
    const float* cp = cpart + (int64_t)j * (j - 1) / 2 * 512 + c;
    float s4[4] = {0.f, 0.f, 0.f, 0.f};
    int i = g;
    for (; i + 12 < j; i += 16) {
        float v[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float* q = cp + (int64_t)(i + 4 * u) * 512;
            v[u][0] = q[0]; v[u][1] = q[128]; v[u][2] = q[256]; v[u][3] = q[384];
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) s4[u] += (v[u][0] + v[u][1]) + (v[u][2] + v[u][3]);
    }
#pragma unroll
    for (int u = 0; u < 3; ++u) {  // up to three tile pairs left
        if (i + 4 * u < j) {
            const float* q = cp + (int64_t)(i + 4 * u) * 512;
            s4[u] += (q[0] + q[128]) + (q[256] + q[384]);
        }
    }
    float s = (s4[0] + s4[1]) + (s4[2] + s4[3]);
    if (g > 0) part[g - 1][c] = s;
    __syncthreads();
    float contrib = 0.f;
    if (g == 0) {
        if (r < M) {
            s = (s + part[0][c]) + (part[1][c] + part[2][c]);
            const int nslots = (T - j + L - 1) / L;
            for (int sl = 0; sl < nslots; ++sl) s += part_s[(int64_t)sl * M + r];
            const float l2 = log2f(s);
            lse_out[r] = l2 * kLn2;
            contrib = (l2 - pos_t[r]) * kLn2;
        }
        contrib = warp_sum(contrib);
        if ((threadIdx.x & 31) == 0) warp_contrib[threadIdx.x >> 5] = contrib;
    }
    __syncthreads();
    if (threadIdx.x == 0)
        block_partials[j] = (warp_contrib[0] + warp_contrib[1]) + (warp_contrib[2] + warp_contrib[3]);
    if (!counter) return;  // (the loss is then summed by a kernel of its own)
    // the loss: the last block to finish (ticket drawn after a fence) adds the per-tile partials in
    // a fixed order -- one launch less on the critical path, still run-to-run deterministic
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) ticket_s = atomicAdd(counter, 1u);
    __syncthreads();
    if (ticket_s != gridDim.x - 1) return;
    __threadfence();
    double acc = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += kSymFinThreads) acc += (double)__ldcg(block_partials + b);
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = kSymFinThreads / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        *loss_out = (float)(red[0] * (double)loss_scale);
        *counter = 0u;
    }
}

// chunk length: long enough to amortise a CTA's prologue (~1.5 tiles), short enough that the
// last CTAs to finish do not leave the other SMs idle for long
int pick_chunk(int T, int slots) {
    int best = 1;
    double best_cost = 1e300;
    for (int L : {4, 6, 8, 12, 16, 20, 24, 32, 48, 64}) {
        if (L > T && L != 4) continue;
        const int items = sym_item_count(T, L);
        const double work = (double)T * (T + 1) / 2 + 1.5 * items;
        const double cost = work / std::min(slots, items) + 0.5 * (std::min(L, T) + 1.5);
        if (cost < best_cost) { best_cost = cost; best = L; }
    }
    return best;
}

template <int D, bool TF32>
cudaError_t launch_sym_t(const CUtensorMap& tmap, int64_t M, float k_scale, const SymPlan& sp, float* part_s,
                         float* cpart, float* pos_t, cudaStream_t stream) {
    using C = Cfg<D, TF32>;
    const int smem = C::kFwdCtasPerSm == 2 ? std::max(C::kFwdSmem, 80 * 1024) : std::max(C::kFwdSmem, 120 * 1024);
    auto kern = tc_fwd_sym_kernel<D, TF32>;
    static unsigned long long attr_done = 0;
    cudaError_t e = set_smem_once(kern, smem, attr_done);
    if (e != cudaSuccess) return e;
    return launch_kernel(kern, dim3((unsigned)sp.items), dim3(kSymThreads), smem, stream, tmap, M, k_scale, sp.T,
                         sp.L, part_s, cpart, pos_t);
}

}  // namespace

bool tc_sym_supported(int64_t M, int64_t D, bool op_bf16, float tau, unsigned flags) {
    if (!op_bf16 || (flags & 4u) || tau < 0.f) return false;  // (tau < 0: not known yet)
    if (kLog2e / tau > kFixedMaxLimit) return false;           // needs the fixed-max form
    if (!(D == 64 || D == 128 || D == 192 || D == 256)) return false;
    return M >= kSymMinRows;
}

SymPlan tc_sym_plan(int64_t M, int64_t D, int num_sms) {
    SymPlan p{};
    p.T = (int)((M + kTile - 1) / kTile);
    const int tile_bytes = (int)(D * 2 * kTile);
    const int stages = tile_bytes <= 16 * 1024 ? 3 : 2;
    const int per_sm = ((1 + stages) * tile_bytes + 2048 <= 112 * 1024) ? 2 : 1;
    p.L = pick_chunk(p.T, num_sms * per_sm);
    p.items = sym_item_count(p.T, p.L);
    p.slots = (p.T + p.L - 1) / p.L;
    p.cpart_floats = (size_t)p.T * (p.T - 1) / 2 * 512;
    return p;
}

cudaError_t tc_launch_fwd_sym(const void* u_all, int64_t M, int64_t D, float tau, const SymPlan& sp,
                              float* part_s, float* cpart, float* pos_t, cudaStream_t stream) {
    CUtensorMap tmap;
    cudaError_t e = cached_tmap_2d(u_all, false, M, D, kTile, &tmap);
    if (e != cudaSuccess) return e;
    const float k_scale = kLog2e / tau;
    switch (D) {
        case 64: e = launch_sym_t<64, false>(tmap, M, k_scale, sp, part_s, cpart, pos_t, stream); break;
        case 128: e = launch_sym_t<128, false>(tmap, M, k_scale, sp, part_s, cpart, pos_t, stream); break;
        case 192: e = launch_sym_t<192, false>(tmap, M, k_scale, sp, part_s, cpart, pos_t, stream); break;
        case 256: e = launch_sym_t<256, false>(tmap, M, k_scale, sp, part_s, cpart, pos_t, stream); break;
        default: return cudaErrorNotSupported;
    }
    return e;
}

cudaError_t tc_launch_fwd_sym_finalize(int64_t M, const SymPlan& sp, const float* part_s, const float* cpart,
                                       const float* pos_t, float* lse_out, float* block_partials,
                                       cudaStream_t stream, unsigned* counter, float* loss_out, float loss_scale) {
    return launch_kernel(fwd_sym_finalize_kernel, dim3((unsigned)sp.T), dim3(kSymFinThreads), 0, stream, M, sp.T,
                         sp.L, part_s, cpart, pos_t, lse_out, block_partials, counter, loss_out, loss_scale);
}

}  // namespace ntxent_b200
