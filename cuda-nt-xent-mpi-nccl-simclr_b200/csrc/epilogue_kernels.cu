// epilogue_kernels.cu -- the small HBM-bound kernels either side of the tile passes:
//   fwd_finalize : split partials -> l_i, positive logit; loss_reduce: deterministic loss sum
//                  (replaces compute_loss_kernel, src/ntxent_kernel.cu:105-134, and its
//                   single-block diagonal walk);
//   bwd_epilogue : split partials -> dU (positive term, 1/(tau M)), optionally the
//                  normalisation Jacobian -> dz and P_{i,pos(i)};
//   finalize_dz  : the Jacobian alone (multi-GPU path, after the reduce-scatter).
#include "common.cuh"

namespace ntxent_b200 {

constexpr int kMaxDChunks = 8;  // D <= 1024: each lane holds 4 floats per 128-wide chunk

// -------------------------------------------------------------- fwd finalize ---
// The positive logit of every row is there already (pos_t, base-2 units: captured by the
// upper-triangle forward kernel from its own MMA, or formed by pos_dot_kernel below), so all that is
// left is one thread per row merging nsplit coalesced partials: ~M * (8 nsplit + 12) bytes.
constexpr int kFinalizeThreads = 128;
__global__ void __launch_bounds__(kFinalizeThreads)
fwd_finalize_rows_kernel(int64_t row_count, int nsplit, const float* __restrict__ part_m,
                         const float* __restrict__ part_s, const float* __restrict__ pos_t,
                         PeerList lse_outs, int64_t lse_offset, float* __restrict__ block_partials) {
    __shared__ float warp_contrib[kFinalizeThreads / 32];
    pdl_trigger();
    pdl_wait();
    const int64_t r = (int64_t)blockIdx.x * kFinalizeThreads + threadIdx.x;
    float contrib = 0.f;
    if (r < row_count) {
        float m = -INFINITY, s = 0.f;
        const float pt = pos_t[r];
        for (int sp0 = 0; sp0 < nsplit; sp0 += 8) {  // eight partial pairs in flight per thread
            float pm[8], ps[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const bool ok = sp0 + j < nsplit;
                pm[j] = ok ? part_m[(int64_t)(sp0 + j) * row_count + r] : -INFINITY;
                ps[j] = ok ? part_s[(int64_t)(sp0 + j) * row_count + r] : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) lse2_merge(m, s, pm[j], ps[j]);
        }
        const float l2 = m + log2f(s);
        const float l = l2 * kLn2;
        for (int i = 0; i < lse_outs.n; ++i) static_cast<float*>(lse_outs.p[i])[lse_offset + r] = l;
        contrib = (l2 - pt) * kLn2;
    }
    contrib = warp_sum(contrib);
    if ((threadIdx.x & 31) == 0) warp_contrib[threadIdx.x >> 5] = contrib;
    __syncthreads();
    if (threadIdx.x == 0) {
        float bsum = 0.f;
#pragma unroll
        for (int w = 0; w < kFinalizeThreads / 32; ++w) bsum += warp_contrib[w];
        block_partials[blockIdx.x] = bsum;
    }
}

// One CTA sums the per-CTA partials in a fixed order (run-to-run deterministic).  A separate
// launch costs less than the fence + same-address atomic of a last-CTA-done scheme here.
__global__ void __launch_bounds__(1024)
loss_reduce_kernel(const float* __restrict__ block_partials, int n, PeerList loss_outs, int loss_index,
                   float loss_scale, PeerSignal signal) {
    __shared__ double red[1024];
    pdl_trigger();
    pdl_wait();
    double acc = 0.0;
    for (int b = threadIdx.x; b < n; b += 1024) acc += (double)block_partials[b];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    // Lane i serves peer i: the partial loss, then (multi-GPU push path) the flag.  This block runs
    // after the finalize kernel has completed, so the signal covers every l_i stored by that
    // kernel; each lane's release store orders its own loss store before its flag, and the peers
    // are written in parallel rather than one NVLink round trip after the other.
    if (threadIdx.x < 32) {
        unsigned e = signal.epoch ? *signal.epoch + 1u : 0u;
        __syncwarp();
        if (threadIdx.x == 0 && signal.epoch) *signal.epoch = e;
        if ((int)threadIdx.x < loss_outs.n) {
            static_cast<float*>(loss_outs.p[threadIdx.x])[loss_index] = (float)(red[0] * (double)loss_scale);
            if (signal.epoch)  // (the C-ABI checks that both lists name the same peers)
                st_release_sys(static_cast<unsigned*>(signal.flags.p[threadIdx.x]) + signal.rank, e);
        }
    }
}

__global__ void __launch_bounds__(32) peer_wait_kernel(PeerWait wait) {
    pdl_trigger();
    pdl_wait();
    if ((int)threadIdx.x < wait.ranks) peer_wait_rank(wait, threadIdx.x, *wait.epoch);
}

cudaError_t launch_peer_wait(const PeerWait& wait, cudaStream_t stream) {
    if (!wait.flags || !wait.epoch || wait.ranks < 1 || wait.ranks > 32) return cudaErrorInvalidValue;
    return launch_kernel(peer_wait_kernel, dim3(1), dim3(32), 0, stream, wait);
}

// Everything after the forward tile pass in ONE kernel, for callers that can give it a ticket
// counter (a device word that is zero before the first call; it leaves every call at zero again):
// 64 rows per block: a thread per row merges the row's split partials (coalesced) and stores l_i (to
// every peer), the warps form the positive logits (fp32 dots of the operand rows, four rows' loads in
// flight), and the last block to finish -- known by the ticket
// it draws after a system-scope fence -- sums the per-block loss partials in a fixed order, stores
// the loss and publishes the signal.  Replaces pos_dot + finalize_rows + loss_reduce: three
// launches on the critical path of a 0.16 ms multi-GPU step.
constexpr int kFusedRowsPerCta = 64;  // 8 warps x 8 rows
template <typename T>
__global__ void __launch_bounds__(256)
fwd_finalize_fused_kernel(const T* __restrict__ U, int64_t M, int D, int64_t row_begin, int64_t row_count,
                          float k_scale, int nsplit, const float* __restrict__ part_m,
                          const float* __restrict__ part_s, const int* __restrict__ pos_index,
                          PeerList lse_outs, int64_t lse_offset, float* __restrict__ block_partials,
                          unsigned* __restrict__ counter, PeerList loss_outs, int loss_index,
                          float loss_scale, PeerSignal signal) {
    __shared__ float l2_s[kFusedRowsPerCta];
    __shared__ float warp_contrib[8];
    __shared__ unsigned ticket_s;
    __shared__ double red[256];
    pdl_trigger();
    pdl_wait();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t rb = (int64_t)blockIdx.x * kFusedRowsPerCta;
    // (1) one thread per row: the split partials, coalesced (consecutive rows are consecutive words
    //     of every split), eight pairs in flight; l_i to every peer's copy
    if (threadIdx.x < kFusedRowsPerCta) {
        const int64_t r = rb + threadIdx.x;
        float l2 = 0.f;
        if (r < row_count) {
            float m = -INFINITY, s = 0.f;
            for (int sp0 = 0; sp0 < nsplit; sp0 += 8) {
                float pm[8], ps[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const bool ok = sp0 + j < nsplit;
                    pm[j] = ok ? part_m[(int64_t)(sp0 + j) * row_count + r] : -INFINITY;
                    ps[j] = ok ? part_s[(int64_t)(sp0 + j) * row_count + r] : 0.f;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) lse2_merge(m, s, pm[j], ps[j]);
            }
            l2 = m + log2f(s);
            const float l = l2 * kLn2;
            for (int i = 0; i < lse_outs.n; ++i) static_cast<float*>(lse_outs.p[i])[lse_offset + r] = l;
        }
        l2_s[threadIdx.x] = l2;
    }
    // (2) every warp: the positive logits of its eight rows (fp32 dot of the operand rows, both in
    //     L2), four rows' loads in flight at a time
    float dots[8];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        float dot[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int64_t r = rb + warp * 8 + h * 4 + q;
            if (r < row_count) {
                const int64_t row = row_begin + r;
                const int64_t pi = pos_index ? (int64_t)pos_index[row] : pair_row(row, M);
                const T* ui = U + row * D;
                const T* up = U + pi * D;
                for (int k = lane * 4; k < D; k += 128) {
                    const float4 a = Elem<T>::ld4(ui + k), b = Elem<T>::ld4(up + k);
                    dot[q] += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) dots[h * 4 + q] = warp_sum(dot[q]);
    }
    __syncthreads();
    float contrib = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q)
        if (rb + warp * 8 + q < row_count) contrib += (l2_s[warp * 8 + q] - dots[q] * k_scale) * kLn2;
    if (lane == 0) warp_contrib[warp] = contrib;
    __syncthreads();
    if (threadIdx.x == 0) {
        float bsum = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) bsum += warp_contrib[w];
        block_partials[blockIdx.x] = bsum;
    }
    // the ticket: every thread's stores (l_i to the peers, the block partial) are fenced at system
    // scope before the block draws it, so whoever draws the last one sees and covers them all
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) ticket_s = atomicAdd(counter, 1u);
    __syncthreads();
    if (ticket_s != gridDim.x - 1) return;
    __threadfence();
    double acc = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += 256) acc += (double)__ldcg(block_partials + b);
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x < 32) {
        unsigned e = signal.epoch ? *signal.epoch + 1u : 0u;
        __syncwarp();
        if (threadIdx.x == 0) {
            if (signal.epoch) *signal.epoch = e;
            *counter = 0u;  // ready for the next call
        }
        if ((int)threadIdx.x < loss_outs.n) {
            static_cast<float*>(loss_outs.p[threadIdx.x])[loss_index] = (float)(red[0] * (double)loss_scale);
            if (signal.epoch)
                st_release_sys(static_cast<unsigned*>(signal.flags.p[threadIdx.x]) + signal.rank, e);
        }
    }
}

cudaError_t launch_fwd_finalize_push(const void* u_all, bool op_bf16, int64_t M, int64_t D,
                                     int64_t row_begin, int64_t row_count, float tau, int nsplit,
                                     const float* part_m, const float* part_s, float* pos_t, const int* pos_index,
                                     const PeerList& lse_outs, int64_t lse_offset, float* block_partials,
                                     const PeerList& loss_outs, int loss_index, float loss_scale,
                                     const PeerSignal& signal, cudaStream_t stream, unsigned* counter) {
    if (counter) {
        const unsigned blocks = (unsigned)((row_count + kFusedRowsPerCta - 1) / kFusedRowsPerCta);
        if (op_bf16)
            return launch_kernel(fwd_finalize_fused_kernel<__nv_bfloat16>, dim3(blocks), dim3(256), 0, stream,
                                 (const __nv_bfloat16*)u_all, M, (int)D, row_begin, row_count, kLog2e / tau, nsplit,
                                 part_m, part_s, pos_index, lse_outs, lse_offset, block_partials, counter, loss_outs,
                                 loss_index, loss_scale, signal);
        return launch_kernel(fwd_finalize_fused_kernel<float>, dim3(blocks), dim3(256), 0, stream,
                             (const float*)u_all, M, (int)D, row_begin, row_count, kLog2e / tau, nsplit, part_m,
                             part_s, pos_index, lse_outs, lse_offset, block_partials, counter, loss_outs, loss_index,
                             loss_scale, signal);
    }
    // the positive logits: k u_i . u_pos(i), fp32 dot products of the operand rows
    cudaError_t e = launch_pos_dot(u_all, op_bf16, M, D, row_begin, row_count, pos_index, kLog2e / tau, pos_t, stream);
    if (e != cudaSuccess) return e;
    const unsigned blocks = (unsigned)((row_count + kFinalizeThreads - 1) / kFinalizeThreads);
    e = launch_kernel(fwd_finalize_rows_kernel, dim3(blocks), dim3(kFinalizeThreads), 0, stream, row_count, nsplit,
                      part_m, part_s, (const float*)pos_t, lse_outs, lse_offset, block_partials);
    if (e != cudaSuccess) return e;
    return launch_loss_reduce(block_partials, (int)blocks, loss_outs, loss_index, loss_scale, signal, stream);
}

cudaError_t launch_loss_reduce(const float* block_partials, int n, const PeerList& loss_outs, int loss_index,
                               float loss_scale, const PeerSignal& signal, cudaStream_t stream) {
    return launch_kernel(loss_reduce_kernel, dim3(1), dim3(1024), 0, stream, block_partials, n, loss_outs,
                         loss_index, loss_scale, signal);
}

// -------------------------------------------------------------- bwd epilogue ---
struct BwdEpiDev {
    const float* acc;
    int nsplit;
    int sk_ctas, sk_col_tiles;  // > 0: the row tile's partial count follows from the tile dealing
    int64_t sk_total;
    int64_t M, a_begin, a_count;
    int D;
    float pos_coeff;
    int64_t pos_lo, pos_hi;
    float scale;
    float* du_out;
    const float* inv_norm;
    const float* grad_out;
    const float* lse_all;
    float inv_tau;
    float* pos_prob_out;
    float* tau_grad_rows_out;
    const int* pos_index;
    const int* inv_ptr;
    const int* inv_idx;
    const float* loss_parts;
    int loss_n;
    float loss_scale;
    float* loss_out;
};

// A warp owns kRows rows of kChunks x 128 columns each and issues the loads of all of them
// (partials, u_pos, z) before the first use: kRows * kChunks 16-byte requests in flight per lane
// and per operand.  (One row per warp, as this kernel used to work, left it latency-bound at
// 19 % of the HBM peak.)
// Rows per warp at D <= 128 and the resident CTAs asked of the compiler (A/B build switches).
// Measured at 32768 x 128 (run Z): 4 rows / 2 CTAs (128 registers) 23.4 us, 2 rows / 4 CTAs 17.6 us.
#ifndef NTX_EPI_ROWS
#define NTX_EPI_ROWS 2
#endif
#ifndef NTX_EPI_MINB
#define NTX_EPI_MINB 4
#endif
template <typename TOp, typename TZ, int kChunks, int kRows>
__global__ void __launch_bounds__(256, kChunks == 1 ? NTX_EPI_MINB : 1)
bwd_epilogue_kernel(BwdEpiDev p, const TOp* __restrict__ U, const TZ* __restrict__ z_rows,
                    TZ* __restrict__ dz_out) {
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int64_t r0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * kRows;
    if (r0 >= p.a_count) return;
    pdl_wait();
    if (p.loss_out && blockIdx.x == 0 && threadIdx.x == 0) {  // the global loss, in rank order
        float acc = 0.f;
        for (int i = 0; i < p.loss_n; ++i) acc += p.loss_parts[i];
        *p.loss_out = acc * p.loss_scale;
    }
    const int D = p.D;
    const bool jac = p.du_out == nullptr;
    // The kRows rows of a warp lie in one 128-row tile (r0 is a multiple of kRows, kRows divides 128):
    // one slot count per warp, worked out first so that the loads of the first two slots go out
    // together with z and u_pos -- one round trip to L2/HBM for the common case instead of three.
    int nsplit = p.nsplit;
    if (p.sk_ctas > 0) {  // CTAs whose tile ranges touch this row tile (tc_kernels.cu, BwdUnit)
        const int64_t first = (r0 / 128) * p.sk_col_tiles;
        if (p.sk_total * p.sk_ctas < (int64_t)0x7fffffff) {  // the usual case: 32-bit divisions
            const unsigned tot = (unsigned)p.sk_total, f32 = (unsigned)first, ctas = (unsigned)p.sk_ctas;
            nsplit = (int)(((f32 + (unsigned)p.sk_col_tiles) * ctas - 1u) / tot - ((f32 + 1u) * ctas - 1u) / tot) + 1;
        } else {
            nsplit = (int)(((first + p.sk_col_tiles) * p.sk_ctas - 1) / p.sk_total -
                           ((first + 1) * p.sk_ctas - 1) / p.sk_total) + 1;
        }
    }
    float4 du[kRows][kChunks], zz[kRows][kChunks], t0[kRows][kChunks], t1[kRows][kChunks];
    int64_t pos[kRows];
    float pc[kRows], inv[kRows];
    auto slot = [&](int sp, int64_t r, int k) {
        return *reinterpret_cast<const float4*>(p.acc + ((int64_t)sp * p.a_count + r) * D + k);
    };
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
        const int64_t r = min(r0 + q, p.a_count - 1);  // rows past the end redo the last row, unstored
        const int64_t row = p.a_begin + r;
        pos[q] = p.pos_index ? (int64_t)p.pos_index[row] : pair_row(row, p.M);
        pc[q] = p.pos_index ? 1.f : ((pos[q] >= p.pos_lo && pos[q] < p.pos_hi) ? p.pos_coeff : 0.f);
        inv[q] = jac ? p.inv_norm[r] : 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            const int k = c * 128 + lane * 4;
            zz[q][c] = make_float4(0.f, 0.f, 0.f, 0.f);
            du[q][c] = zz[q][c];
            t0[q][c] = zz[q][c];
            t1[q][c] = zz[q][c];
            if (k < D) {
                if (nsplit > 0) t0[q][c] = slot(0, r, k);
                if (nsplit > 1) t1[q][c] = slot(1, r, k);
                if (jac) zz[q][c] = Elem<TZ>::ld4(z_rows + r * D + k);
                if (pc[q] != 0.f) {
                    const float4 u = Elem<TOp>::ld4(U + pos[q] * D + k);
                    du[q][c] = make_float4(-pc[q] * u.x, -pc[q] * u.y, -pc[q] * u.z, -pc[q] * u.w);
                }
            }
        }
    }
    // positives from an index vector: the rows that name this one as THEIR positive, in list order
    if (p.pos_index) {
#pragma unroll
        for (int q = 0; q < kRows; ++q) {
            const int64_t row = p.a_begin + min(r0 + q, p.a_count - 1);
            for (int e = p.inv_ptr[row]; e < p.inv_ptr[row + 1]; ++e) {
                const TOp* ub = U + (int64_t)p.inv_idx[e] * D;
#pragma unroll
                for (int c = 0; c < kChunks; ++c) {
                    const int k = c * 128 + lane * 4;
                    if (k < D) {
                        const float4 u = Elem<TOp>::ld4(ub + k);
                        du[q][c].x -= u.x; du[q][c].y -= u.y; du[q][c].z -= u.z; du[q][c].w -= u.w;
                    }
                }
            }
        }
    }
    // the partials of the row tile, added in slot order (deterministic); slots past the second (small
    // row blocks against many columns) a slot at a time, all rows' loads before the adds
#pragma unroll
    for (int q = 0; q < kRows; ++q)
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            t0[q][c].x += t1[q][c].x; t0[q][c].y += t1[q][c].y;
            t0[q][c].z += t1[q][c].z; t0[q][c].w += t1[q][c].w;
        }
    for (int sp = 2; sp < nsplit; ++sp) {
#pragma unroll
        for (int q = 0; q < kRows; ++q) {
            const int64_t r = min(r0 + q, p.a_count - 1);
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const int k = c * 128 + lane * 4;
                t1[q][c] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (k < D) t1[q][c] = slot(sp, r, k);
            }
        }
#pragma unroll
        for (int q = 0; q < kRows; ++q)
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                t0[q][c].x += t1[q][c].x; t0[q][c].y += t1[q][c].y;
                t0[q][c].z += t1[q][c].z; t0[q][c].w += t1[q][c].w;
            }
    }
#pragma unroll
    for (int q = 0; q < kRows; ++q)
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            du[q][c].x = (du[q][c].x + t0[q][c].x) * p.scale;
            du[q][c].y = (du[q][c].y + t0[q][c].y) * p.scale;
            du[q][c].z = (du[q][c].z + t0[q][c].z) * p.scale;
            du[q][c].w = (du[q][c].w + t0[q][c].w) * p.scale;
        }
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
        const int64_t r = r0 + q;
        if (r >= p.a_count) break;
        const int64_t row = p.a_begin + r;
        if (p.pos_prob_out) {
            const TOp* ui = U + row * D;
            const TOp* up = U + pos[q] * D;
            float dot = 0.f;
            for (int k = lane * 4; k < D; k += 128) {
                const float4 a = Elem<TOp>::ld4(ui + k), b = Elem<TOp>::ld4(up + k);
                dot += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
            }
            dot = warp_sum(dot);
            if (lane == 0) p.pos_prob_out[r] = expf(dot * p.inv_tau - p.lse_all[row]);
        }
        if (!jac) {
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const int k = c * 128 + lane * 4;
                if (k < D) *reinterpret_cast<float4*>(p.du_out + r * D + k) = du[q][c];
            }
            continue;
        }
        // normalisation Jacobian with the exact fp32 direction u = z * inv_norm
        const float g = p.grad_out ? *p.grad_out : 1.0f;
        float dot = 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            float4& u = zz[q][c];
            u.x *= inv[q]; u.y *= inv[q]; u.z *= inv[q]; u.w *= inv[q];
            dot += du[q][c].x * u.x + du[q][c].y * u.y + du[q][c].z * u.z + du[q][c].w * u.w;
        }
        dot = warp_sum(dot);
        // sum_a dU_a . u_a = -2 tau dL/dtau (each logit S_ab = u_a.u_b / tau appears once as row a's
        // and once as row b's): the temperature gradient falls out of the Jacobian's dot product
        if (p.tau_grad_rows_out && lane == 0) p.tau_grad_rows_out[r] = -0.5f * p.inv_tau * dot;
        const float gi = g * inv[q];
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            const int k = c * 128 + lane * 4;
            if (k >= D) break;
            float4 o;
            o.x = gi * (du[q][c].x - dot * zz[q][c].x);
            o.y = gi * (du[q][c].y - dot * zz[q][c].y);
            o.z = gi * (du[q][c].z - dot * zz[q][c].z);
            o.w = gi * (du[q][c].w - dot * zz[q][c].w);
            Elem<TZ>::st4(dz_out + r * D + k, o);
        }
    }
}

template <typename TOp, typename TZ>
static cudaError_t launch_bwd_epilogue_t(const BwdEpiDev& p, const BwdEpilogueArgs& a, cudaStream_t stream) {
    const int chunks = (p.D + 127) / 128;
    dim3 block(256);
#define NTX_EPI(C, R)                                                                                   \
    launch_kernel(bwd_epilogue_kernel<TOp, TZ, C, R>, dim3((unsigned)((a.a_count + 8 * R - 1) / (8 * R))), \
                  block, 0, stream, p, (const TOp*)a.u_all, (const TZ*)a.z_rows, (TZ*)a.dz_out)
    if (chunks <= 1) return NTX_EPI(1, NTX_EPI_ROWS);
    if (chunks <= 2) return NTX_EPI(2, 2);
    if (chunks <= 4) return NTX_EPI(4, 1);
    return NTX_EPI(8, 1);
#undef NTX_EPI
}

cudaError_t launch_bwd_epilogue(const BwdEpilogueArgs& a, cudaStream_t stream) {
    if (a.a_count == 0) return cudaSuccess;
    BwdEpiDev p;
    p.sk_ctas = a.sk_ctas; p.sk_col_tiles = a.sk_col_tiles;
    p.sk_total = (int64_t)((a.a_count + 127) / 128) * a.sk_col_tiles;
    p.acc = a.acc; p.nsplit = a.nsplit; p.M = a.M; p.a_begin = a.a_begin; p.a_count = a.a_count;
    p.D = (int)a.D; p.pos_coeff = a.pos_coeff; p.pos_lo = a.pos_lo; p.pos_hi = a.pos_hi;
    p.scale = a.scale; p.du_out = a.du_out; p.inv_norm = a.inv_norm; p.grad_out = a.grad_out;
    p.lse_all = a.lse_all; p.inv_tau = a.inv_tau; p.pos_prob_out = a.pos_prob_out;
    p.tau_grad_rows_out = a.du_out ? nullptr : a.tau_grad_rows_out;
    p.pos_index = a.pos_index; p.inv_ptr = a.inv_ptr; p.inv_idx = a.inv_idx;
    p.loss_parts = a.loss_parts; p.loss_n = a.loss_n; p.loss_scale = a.loss_scale; p.loss_out = a.loss_out;
    if (a.op_bf16)
        return a.z_bf16 ? launch_bwd_epilogue_t<__nv_bfloat16, __nv_bfloat16>(p, a, stream)
                        : launch_bwd_epilogue_t<__nv_bfloat16, float>(p, a, stream);
    return a.z_bf16 ? launch_bwd_epilogue_t<float, __nv_bfloat16>(p, a, stream)
                    : launch_bwd_epilogue_t<float, float>(p, a, stream);
}

// ------------------------------------------------------------------ pos dot ---
// pos_t[r] = k u_i . u_pos(i) for rows i = row_begin + r: the positive (i + M/2) mod M, or
// pos_index[i].  One warp per four rows, all loads issued first (both rows of a pair live in L2).
template <typename T>
__global__ void __launch_bounds__(256)
pos_dot_kernel(const T* __restrict__ U, int64_t M, int D, int64_t row_begin, int64_t row_count,
               const int* __restrict__ pos_index, float k_scale, float* __restrict__ pos_t) {
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int64_t r0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 4;
    if (r0 >= row_count) return;
    pdl_wait();
    float dot[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int64_t i = row_begin + min(r0 + q, row_count - 1);
        const int64_t pi = pos_index ? (int64_t)pos_index[i] : pair_row(i, M);
        const T* ui = U + i * D;
        const T* up = U + pi * D;
        for (int k = lane * 4; k < D; k += 128) {
            const float4 a = Elem<T>::ld4(ui + k), b = Elem<T>::ld4(up + k);
            dot[q] += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float d = warp_sum(dot[q]);
        if (lane == 0 && r0 + q < row_count) pos_t[r0 + q] = d * k_scale;
    }
}

cudaError_t launch_pos_dot(const void* u_all, bool op_bf16, int64_t M, int64_t D, int64_t row_begin,
                           int64_t row_count, const int* pos_index, float k_scale, float* pos_t,
                           cudaStream_t stream) {
    dim3 grid((unsigned)((row_count + 31) / 32)), block(256);
    if (op_bf16)
        return launch_kernel(pos_dot_kernel<__nv_bfloat16>, grid, block, 0, stream, (const __nv_bfloat16*)u_all, M,
                             (int)D, row_begin, row_count, pos_index, k_scale, pos_t);
    return launch_kernel(pos_dot_kernel<float>, grid, block, 0, stream, (const float*)u_all, M, (int)D, row_begin,
                         row_count, pos_index, k_scale, pos_t);
}

// --------------------------------------------------------------- finalize dz ---
template <typename TZ>
__global__ void __launch_bounds__(256)
finalize_dz_kernel(const float* __restrict__ du_rows, const TZ* __restrict__ z_rows,
                   const float* __restrict__ inv_norm, const float* __restrict__ grad_out,
                   int64_t rows, int D, TZ* __restrict__ dz_out) {
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    pdl_wait();
    const float inv = inv_norm[r];
    const float g = grad_out ? *grad_out : 1.0f;
    float dot = 0.f;
    for (int k = lane * 4; k < D; k += 128) {
        const float4 d = *reinterpret_cast<const float4*>(du_rows + r * D + k);
        const float4 zz = Elem<TZ>::ld4(z_rows + r * D + k);
        dot += (d.x * zz.x + d.y * zz.y + d.z * zz.z + d.w * zz.w) * inv;
    }
    dot = warp_sum(dot);
    const float gi = g * inv;
    for (int k = lane * 4; k < D; k += 128) {
        const float4 d = *reinterpret_cast<const float4*>(du_rows + r * D + k);
        const float4 zz = Elem<TZ>::ld4(z_rows + r * D + k);
        float4 o;
        o.x = gi * (d.x - dot * zz.x * inv);
        o.y = gi * (d.y - dot * zz.y * inv);
        o.z = gi * (d.z - dot * zz.z * inv);
        o.w = gi * (d.w - dot * zz.w * inv);
        Elem<TZ>::st4(dz_out + r * D + k, o);
    }
}

cudaError_t launch_finalize_dz(const float* du_rows, const void* z_rows, bool z_bf16,
                               const float* inv_norm, const float* grad_out, int64_t rows, int64_t D,
                               void* dz_out, cudaStream_t stream) {
    if (rows == 0) return cudaSuccess;
    dim3 grid((unsigned)((rows + 7) / 8)), block(256);
    if (z_bf16)
        return launch_kernel(finalize_dz_kernel<__nv_bfloat16>, grid, block, 0, stream, du_rows,
                             (const __nv_bfloat16*)z_rows, inv_norm, grad_out, rows, (int)D,
                             (__nv_bfloat16*)dz_out);
    return launch_kernel(finalize_dz_kernel<float>, grid, block, 0, stream, du_rows, (const float*)z_rows,
                         inv_norm, grad_out, rows, (int)D, (float*)dz_out);
}

}  // namespace ntxent_b200
