// simt_kernels.cu -- row normalisation and the fp32-FMA tile kernels.
//
// These are the exact-fp32 members of the engine: the path taken for fp32 operands (the
// north_star's 1e-5 / 1e-4 tolerance case), for embedding widths the tcgen05 tiles do not
// cover (D not a multiple of 64, or D > 256) and when NTXENT_FLAG_FORCE_SIMT is set.  They
// share everything around them (normalise -> tile pass -> finalize/epilogue, the split plan,
// the workspace) with the tcgen05 kernels in tc_kernels.cu.
//
// They replace, for this path, the reference's materialised pipeline
// cublasSgemm -> row_max_kernel -> softmax_kernel -> compute_loss_kernel
// (src/ntxent_kernel.cu:162-199) and its backward (src/ntxent_kernel.cu:217-236); the
// [2B,2B] logits/softmax tensors are never formed.
#include <algorithm>

#include "common.cuh"

namespace ntxent_b200 {

// ------------------------------------------------------------------ normalise ---
// u = z / max(|z|, eps), computed as (z/amax) / |z/amax| so that rows scaled by 1e-5 .. 1e5
// (python/test.py:64) neither underflow nor overflow.  One pass: a warp keeps kRows whole rows
// in registers (kChunks x 128 columns each), with every load issued before the first use, so
// that each lane has kRows * kChunks 16-byte (fp32) or 8-byte (bf16) requests in flight -- the
// one-row-per-warp, three-pass version this replaces ran at 21 % of the HBM peak.
template <typename TIn, typename TOut, int kChunks, int kRows>
__global__ void __launch_bounds__(256) normalize_kernel(const TIn* __restrict__ z, int64_t rows,
                                                        int D, PeerList outs, int64_t row_offset,
                                                        float* __restrict__ inv_norm, unsigned* zero_words, int n_zero,
                                                        PeerSignal signal) {
    __shared__ unsigned ticket_s;
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const int64_t row0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * kRows;
    pdl_wait();  // the previous step's epilogue may still be reading the U this kernel overwrites
    if (zero_words && blockIdx.x == 0)
        for (int i = threadIdx.x; i < n_zero; i += 256) zero_words[i] = 0u;
    if (row0 < rows) {
    float4 v[kRows][kChunks];
#pragma unroll
    for (int q = 0; q < kRows; ++q)
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            const int k = c * 128 + lane * 4;
            v[q][c] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row0 + q < rows && k < D) v[q][c] = Elem<TIn>::ld4(z + (row0 + q) * D + k);
        }
#pragma unroll
    for (int q = 0; q < kRows; ++q) {
        float amax = 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c)
            amax = fmaxf(amax, fmaxf(fmaxf(fabsf(v[q][c].x), fabsf(v[q][c].y)),
                                     fmaxf(fabsf(v[q][c].z), fabsf(v[q][c].w))));
        amax = warp_max(amax);
        const float sc = (amax > 0.f && amax < INFINITY) ? 1.0f / amax : 1.0f;
        float ss = 0.f;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            const float4 t = make_float4(v[q][c].x * sc, v[q][c].y * sc, v[q][c].z * sc, v[q][c].w * sc);
            ss += t.x * t.x + t.y * t.y + t.z * t.z + t.w * t.w;
        }
        ss = warp_sum(ss);
        const float nrm = sqrtf(ss) / sc;  // |z|
        const float inv = 1.0f / fmaxf(nrm, kNormEps);
        const int64_t row = row0 + q;
        if (row >= rows) break;
        if (lane == 0 && inv_norm) inv_norm[row] = inv;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
            const int k = c * 128 + lane * 4;
            if (k >= D) break;
            const float4 t = make_float4(v[q][c].x * inv, v[q][c].y * inv, v[q][c].z * inv, v[q][c].w * inv);
            // one local store, or the same row pushed into every rank's gathered copy (peer stores)
            for (int i = 0; i < outs.n; ++i)
                Elem<TOut>::st4(static_cast<TOut*>(outs.p[i]) + (row_offset + row) * D + k, t);
        }
    }
    }
    // Push path with a ticket word: the block that finishes last publishes "this rank's rows have
    // landed" itself (every block fences its peer stores at system scope before it draws its ticket),
    // which saves the one-block signal kernel and its launch boundary.
    if (signal.ticket) {
        __threadfence_system();
        __syncthreads();
        if (threadIdx.x == 0) ticket_s = atomicAdd(signal.ticket, 1u);
        __syncthreads();
        if (ticket_s == gridDim.x - 1 && threadIdx.x < 32) {
            unsigned e = 0;
            if (threadIdx.x == 0) {
                e = *signal.epoch + 1u;
                *signal.epoch = e;
                *signal.ticket = 0u;
            }
            e = __shfl_sync(0xffffffffu, e, 0);
            if ((int)threadIdx.x < signal.flags.n)
                st_release_sys(static_cast<unsigned*>(signal.flags.p[threadIdx.x]) + signal.rank, e);
        }
    }
}

// Publishes "the kernel before me on this stream has finished its peer stores": a grid's stores
// are complete when the next kernel of the stream starts, so one thread bumping the epoch and
// writing it to every rank's slot is all it takes -- measured far cheaper than ending each of
// the ~2000 normalize blocks with a system-scope fence (+11 us at 16384 x 128 on two GPUs).
__global__ void __launch_bounds__(32) peer_signal_kernel(PeerSignal s) {
    pdl_trigger();
    pdl_wait();  // (this is the "kernel before me has finished" the comment above relies on)
    unsigned e = 0;
    if (threadIdx.x == 0) {
        e = *s.epoch + 1u;
        *s.epoch = e;
    }
    e = __shfl_sync(0xffffffffu, e, 0);
    if ((int)threadIdx.x < s.flags.n) st_release_sys(static_cast<unsigned*>(s.flags.p[threadIdx.x]) + s.rank, e);
}

template <typename TIn, typename TOut>
static cudaError_t launch_normalize_t(const void* z, int64_t rows, int64_t D, const PeerList& outs,
                                      int64_t row_offset, float* inv_norm, cudaStream_t stream,
                                      unsigned* zero_words, int n_zero, const PeerSignal& signal) {
    const int chunks = (int)((D + 127) / 128);
    dim3 block(256);
#define NTX_NORM(C, R)                                                                           \
    launch_kernel(normalize_kernel<TIn, TOut, C, R>, dim3((unsigned)((rows + 8 * R - 1) / (8 * R))), \
                  block, 0, stream, (const TIn*)z, rows, (int)D, outs, row_offset, inv_norm, zero_words, n_zero, signal)
    if (chunks <= 1) return NTX_NORM(1, 4);
    if (chunks <= 2) return NTX_NORM(2, 4);
    if (chunks <= 4) return NTX_NORM(4, 2);
    return NTX_NORM(8, 1);
#undef NTX_NORM
}

cudaError_t launch_normalize_push(const void* z, bool z_bf16, int64_t rows, int64_t D,
                                  const PeerList& outs, int64_t row_offset, bool op_bf16,
                                  float* inv_norm, const PeerSignal& signal, cudaStream_t stream,
                                  unsigned* zero_words, int n_zero) {
    if (rows == 0) return signal.epoch ? cudaErrorInvalidValue : cudaSuccess;
    // with a ticket word the kernel publishes the signal itself; otherwise a one-block kernel after it
    PeerSignal in_kernel = no_signal();
    if (signal.epoch && signal.ticket) in_kernel = signal;
    cudaError_t e;
    if (z_bf16 && op_bf16)
        e = launch_normalize_t<__nv_bfloat16, __nv_bfloat16>(z, rows, D, outs, row_offset, inv_norm, stream, zero_words, n_zero, in_kernel);
    else if (z_bf16)
        e = launch_normalize_t<__nv_bfloat16, float>(z, rows, D, outs, row_offset, inv_norm, stream, zero_words, n_zero, in_kernel);
    else if (op_bf16)
        e = launch_normalize_t<float, __nv_bfloat16>(z, rows, D, outs, row_offset, inv_norm, stream, zero_words, n_zero, in_kernel);
    else
        e = launch_normalize_t<float, float>(z, rows, D, outs, row_offset, inv_norm, stream, zero_words, n_zero, in_kernel);
    if (e != cudaSuccess || !signal.epoch || signal.ticket) return e;
    return launch_kernel(peer_signal_kernel, dim3(1), dim3(32), 0, stream, signal);
}

cudaError_t launch_normalize(const void* z, bool z_bf16, int64_t rows, int64_t D, void* u_out,
                             bool op_bf16, float* inv_norm, cudaStream_t stream, unsigned* zero_words, int n_zero) {
    return launch_normalize_push(z, z_bf16, rows, D, single_peer(u_out), 0, op_bf16, inv_norm, no_signal(),
                                 stream, zero_words, n_zero);
}

// ------------------------------------------------------------- 64x64 S tile ---
constexpr int kTM = 64, kTN = 64, kTK = 16, kPad = 4;

// acc[a][b] = u_{a0+ty*4+a} . u_{b0+tx*4+b} (fp32), rows past M read as zero.
template <typename T>
__device__ __forceinline__ void simt_s_tile(const T* __restrict__ U, int64_t M, int D, int64_t a0,
                                            int64_t b0, float (&acc)[4][4],
                                            float (*As)[kTM + kPad], float (*Bs)[kTN + kPad]) {
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int lr = tid >> 2, lk = (tid & 3) * 4;
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
    for (int k0 = 0; k0 < D; k0 += kTK) {
        float4 va = make_float4(0.f, 0.f, 0.f, 0.f), vb = va;
        if (k0 + lk < D) {
            if (a0 + lr < M) va = Elem<T>::ld4(U + (a0 + lr) * D + k0 + lk);
            if (b0 + lr < M) vb = Elem<T>::ld4(U + (b0 + lr) * D + k0 + lk);
        }
        __syncthreads();
        As[lk + 0][lr] = va.x; As[lk + 1][lr] = va.y; As[lk + 2][lr] = va.z; As[lk + 3][lr] = va.w;
        Bs[lk + 0][lr] = vb.x; Bs[lk + 1][lr] = vb.y; Bs[lk + 2][lr] = vb.z; Bs[lk + 3][lr] = vb.w;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kTK; ++k) {
            const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a4.x, a4.y, a4.z, a4.w}, bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
        }
    }
}

// ---------------------------------------------------------------- forward ---
// grid (row tiles, column splits).  Each CTA walks its column range and keeps a running
// base-2 (max, sum) per row; partials go to part_m/part_s[split][row_count].
template <typename T>
__global__ void __launch_bounds__(256)
simt_fwd_kernel(const T* __restrict__ U, int64_t M, int D, int64_t row_begin, int64_t row_count,
                float scale_log2, int64_t cols_per_split, int64_t cross_n, float* __restrict__ part_m,
                float* __restrict__ part_s, PeerWait u_ready) {
    __shared__ __align__(16) float As[kTK][kTM + kPad];
    __shared__ __align__(16) float Bs[kTK][kTN + kPad];
    pdl_trigger();
    pdl_wait();
    peer_wait_all_block(u_ready);  // multi-GPU push path: every rank's rows of U have landed
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t a0 = row_begin + (int64_t)blockIdx.x * kTM;
    const int64_t b_lo = (int64_t)blockIdx.y * cols_per_split;
    const int64_t b_hi = min(M, b_lo + cols_per_split);
    float m[4], s[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) { m[a] = -INFINITY; s[a] = 0.f; }
    float acc[4][4];
    for (int64_t b0 = b_lo; b0 < b_hi; b0 += kTN) {
        simt_s_tile<T>(U, M, D, a0, b0, acc, As, Bs);
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int64_t row = a0 + ty * 4 + a;
            float t[4], tmax = -INFINITY;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int64_t col = b0 + tx * 4 + b;
                // cross_n > 0: only the other view's columns count (two-tower form)
                const bool ok = (col < b_hi) && (col != row) && !(cross_n > 0 && ((col >= cross_n) == (row >= cross_n)));
                t[b] = ok ? acc[a][b] * scale_log2 : -INFINITY;
                tmax = fmaxf(tmax, t[b]);
            }
            const float mn = fmaxf(m[a], tmax);
            if (mn != -INFINITY) {
                float add = 0.f;
#pragma unroll
                for (int b = 0; b < 4; ++b) add += exp2f(t[b] - mn);
                s[a] = s[a] * exp2f(m[a] - mn) + add;
                m[a] = mn;
            }
        }
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, m[a], o);
            const float s2 = __shfl_xor_sync(0xffffffffu, s[a], o);
            lse2_merge(m[a], s[a], m2, s2);
        }
        const int64_t r = (int64_t)blockIdx.x * kTM + ty * 4 + a;
        if (tx == 0 && r < row_count) {
            part_m[(int64_t)blockIdx.y * row_count + r] = m[a];
            part_s[(int64_t)blockIdx.y * row_count + r] = s[a];
        }
    }
}

cudaError_t launch_simt_fwd(const void* u_all, bool op_bf16, int64_t M, int64_t D,
                            int64_t row_begin, int64_t row_count, float tau, const SplitPlan& plan,
                            float* part_m, float* part_s, const PeerWait& u_ready,
                            cudaStream_t stream, int64_t cross_n) {
    dim3 grid((unsigned)plan.row_tiles, (unsigned)plan.nsplit), block(256);
    const float scale_log2 = kLog2e / tau;
    if (op_bf16)
        return launch_kernel(simt_fwd_kernel<__nv_bfloat16>, grid, block, 0, stream,
                             (const __nv_bfloat16*)u_all, M, (int)D, row_begin, row_count, scale_log2,
                             plan.cols_per_split, cross_n, part_m, part_s, u_ready);
    return launch_kernel(simt_fwd_kernel<float>, grid, block, 0, stream, (const float*)u_all, M, (int)D,
                         row_begin, row_count, scale_log2, plan.cols_per_split, cross_n, part_m, part_s, u_ready);
}

// --------------------------------------------------------------- backward ---
// grid (A row tiles, B column splits, D chunks of 128).  Per column tile: recompute S,
// form the weights W (mode), stage them in shared memory and accumulate W . U_b into a
// 64 x 128 register tile.  Output: acc[split][a_count][D] raw sums (scaled in the epilogue).
constexpr int kDC = 128;

template <typename T, int MODE>
__global__ void __launch_bounds__(256)
simt_bwd_kernel(const T* __restrict__ U, const float* lse, int64_t M, int D,
                int64_t a_begin, int64_t a_count, int64_t b_begin, int64_t b_count,
                float scale_log2, int64_t cols_per_split, int64_t cross_n, float* __restrict__ acc_out,
                PeerWait lse_ready) {
    pdl_trigger();
    pdl_wait();
    peer_wait_all_block(lse_ready);  // multi-GPU push path: every rank's l has landed
    __shared__ __align__(16) float As[kTK][kTM + kPad];
    __shared__ __align__(16) float Bs[kTK][kTN + kPad];
    __shared__ float Ws[kTM][kTN + 1];
    __shared__ __align__(16) float Vs[kTK][kDC + kPad];
    __shared__ float lcol[kTN];
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t a0 = a_begin + (int64_t)blockIdx.x * kTM;
    const int64_t a_end = a_begin + a_count;
    const int64_t b_lo = b_begin + (int64_t)blockIdx.y * cols_per_split;
    const int64_t b_hi = min(b_begin + b_count, b_lo + cols_per_split);
    const int dc0 = blockIdx.z * kDC;
    float lrow[4];
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t row = a0 + ty * 4 + a;
        lrow[a] = (MODE != kCol && row < a_end) ? lse[row] * kLog2e : 0.f;
    }
    float o[4][8];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int c = 0; c < 8; ++c) o[a][c] = 0.f;
    float acc[4][4];
    for (int64_t b0 = b_lo; b0 < b_hi; b0 += kTN) {
        simt_s_tile<T>(U, M, D, a0, b0, acc, As, Bs);
        if (MODE != kRow && tid < kTN) lcol[tid] = (b0 + tid < b_hi) ? lse[b0 + tid] * kLog2e : 0.f;
        __syncthreads();
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int64_t row = a0 + ty * 4 + a;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int64_t col = b0 + tx * 4 + b;
                const float t = acc[a][b] * scale_log2;
                float w = 0.f;
                if (col < b_hi && col != row && row < a_end &&
                    !(cross_n > 0 && ((col >= cross_n) == (row >= cross_n)))) {
                    if (MODE != kCol) w += exp2f(t - lrow[a]);
                    if (MODE != kRow) w += exp2f(t - lcol[tx * 4 + b]);
                }
                Ws[ty * 4 + a][tx * 4 + b] = w;
            }
        }
        for (int j0 = 0; j0 < kTN; j0 += kTK) {
            // stage U[b0+j0 .. +15][dc0 .. dc0+127]
            {
                const int jr = tid >> 4, c = (tid & 15) * 8;
                const int64_t brow = b0 + j0 + jr;
                float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
                if (brow < b_hi && dc0 + c < D) {
                    v0 = Elem<T>::ld4(U + brow * D + dc0 + c);
                    v1 = Elem<T>::ld4(U + brow * D + dc0 + c + 4);
                }
                __syncthreads();  // W stores (first round) / previous Vs reads complete
                *reinterpret_cast<float4*>(&Vs[jr][c]) = v0;
                *reinterpret_cast<float4*>(&Vs[jr][c + 4]) = v1;
                __syncthreads();
            }
#pragma unroll
            for (int j = 0; j < kTK; ++j) {
                const float4 v0 = *reinterpret_cast<const float4*>(&Vs[j][tx * 4]);
                const float4 v1 = *reinterpret_cast<const float4*>(&Vs[j][64 + tx * 4]);
                const float vv[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    const float w = Ws[ty * 4 + a][j0 + j];
#pragma unroll
                    for (int c = 0; c < 8; ++c) o[a][c] = fmaf(w, vv[c], o[a][c]);
                }
            }
        }
        __syncthreads();  // Ws / lcol reuse by the next column tile
    }
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t r = (int64_t)blockIdx.x * kTM + ty * 4 + a;
        if (r >= a_count) continue;
        float* dst = acc_out + ((int64_t)blockIdx.y * a_count + r) * D + dc0;
        if (dc0 + tx * 4 < D)
            *reinterpret_cast<float4*>(dst + tx * 4) = make_float4(o[a][0], o[a][1], o[a][2], o[a][3]);
        if (dc0 + 64 + tx * 4 < D)
            *reinterpret_cast<float4*>(dst + 64 + tx * 4) =
                make_float4(o[a][4], o[a][5], o[a][6], o[a][7]);
    }
}

template <typename T>
static cudaError_t launch_simt_bwd_t(const T* U, const float* lse, int64_t M, int64_t D,
                                     int64_t a_begin, int64_t a_count, int64_t b_begin,
                                     int64_t b_count, float tau, int mode, const SplitPlan& plan,
                                     float* acc, const PeerWait& lse_ready, cudaStream_t stream, int64_t cross_n) {
    dim3 grid((unsigned)plan.row_tiles, (unsigned)plan.nsplit, (unsigned)((D + kDC - 1) / kDC));
    dim3 block(256);
    const float sl2 = kLog2e / tau;
#define NTX_SIMT_BWD(MODE)                                                                            \
    launch_kernel(simt_bwd_kernel<T, MODE>, grid, block, 0, stream, U, lse, M, (int)D, a_begin, a_count, \
                  b_begin, b_count, sl2, plan.cols_per_split, cross_n, acc, lse_ready)
    switch (mode) {
        case kSym: return NTX_SIMT_BWD(kSym);
        case kRow: return NTX_SIMT_BWD(kRow);
        default: return NTX_SIMT_BWD(kCol);
    }
#undef NTX_SIMT_BWD
}

cudaError_t launch_simt_bwd(const void* u_all, bool op_bf16, const float* lse_all, int64_t M,
                            int64_t D, int64_t a_begin, int64_t a_count, int64_t b_begin,
                            int64_t b_count, float tau, int mode, const SplitPlan& plan, float* acc,
                            const PeerWait& lse_ready, cudaStream_t stream, int64_t cross_n) {
    if (op_bf16)
        return launch_simt_bwd_t<__nv_bfloat16>((const __nv_bfloat16*)u_all, lse_all, M, D, a_begin,
                                                a_count, b_begin, b_count, tau, mode, plan, acc,
                                                lse_ready, stream, cross_n);
    return launch_simt_bwd_t<float>((const float*)u_all, lse_all, M, D, a_begin, a_count, b_begin,
                                    b_count, tau, mode, plan, acc, lse_ready, stream, cross_n);
}

}  // namespace ntxent_b200
