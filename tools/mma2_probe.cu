// CTA-pair tcgen05.mma (cta_group::2, M = 256): (1) does it compute what the backward kernel would
// need -- which half of B each CTA supplies, where D lands, A from each CTA's own TMEM, B read
// MN-major, the multicast commit -- checked against integer-valued operands, and (2) how many
// clocks one 256xNx16 MMA takes compared with the single-CTA 128xNx16 of tools/mma_rate_probe.cu.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o mma2_probe tools/mma2_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "../cuda-nt-xent-mpi-nccl-simclr_b200/csrc/tc_ptx.cuh"
using namespace ntxent_b200::ptx;

namespace {
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* smem_out) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(smem_out)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_commit2(uint64_t* bar) {  // one arrival on `bar` in BOTH CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                     smem_u32(bar)), "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ void mma2_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, bool acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(a), "l"(b), "r"(idesc),
                 "r"((uint32_t)acc)
                 : "memory");
}
__device__ __forceinline__ void mma2_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, bool acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d), "r"(a), "l"(b), "r"(idesc),
                 "r"((uint32_t)acc)
                 : "memory");
}
// integer-valued operands: exact in bf16, sums exact in fp32
__host__ __device__ inline int a_val(int row, int k) { return (row * 3 + k * 5) % 7 - 3; }
__host__ __device__ inline int b_val(int n, int k) { return (n * 2 + k * 3) % 5 - 2; }
__device__ __forceinline__ uint32_t sw128(int row, int col) {  // byte offset of bf16 (row, col), col < 64
    return row * 128 + ((((col >> 3) ^ (row & 7))) << 4) + (col & 7) * 2;
}
constexpr int kK = 64;  // one 128-byte atom along K: four K = 16 MMAs
// shared-memory map (bytes): A K-major [128 x 64] | B K-major [N/2 x 64] (<= 16 KB) | B MN-major, atoms of [64 k x 64 n]
constexpr int kOffA = 0, kOffBk = 16384, kOffBmn = 32768, kSmem = 65536;
}  // namespace

// form: 0 SS (A, B K-major), 1 TS (A in TMEM, B K-major), 2 TS (B MN-major)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
probe2(int form, int n_dim, int batches, int batch, int validate, unsigned long long* clk, int* bad) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint32_t slot;
    __shared__ uint64_t bar[2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const int pair = blockIdx.x >> 1;
    const int half = n_dim / 2;  // B columns this CTA supplies
    if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_mbar_init(); }
    for (int i = threadIdx.x; i < kSmem / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    __syncthreads();
    if (validate) {
        for (int i = threadIdx.x; i < 128 * kK; i += blockDim.x) {
            const int row = i / kK, k = i % kK;
            *reinterpret_cast<__nv_bfloat16*>(smem + kOffA + sw128(row, k)) = __float2bfloat16((float)a_val(128 * rank + row, k));
        }
        for (int i = threadIdx.x; i < half * kK; i += blockDim.x) {
            const int j = i / kK, k = i % kK, n = half * rank + j;
            *reinterpret_cast<__nv_bfloat16*>(smem + kOffBk + sw128(j, k)) = __float2bfloat16((float)b_val(n, k));
            // MN-major: row = k, 64 n per atom, atoms kK*128 bytes apart
            *reinterpret_cast<__nv_bfloat16*>(smem + kOffBmn + (j >> 6) * (kK * 128) + sw128(k, j & 63)) =
                __float2bfloat16((float)b_val(n, k));
        }
    }
    if (warp == 0) tmem_alloc2(&slot);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tb = slot;
    const uint32_t lane_base = (uint32_t)(warp * 32) << 16;
    {   // A into TMEM columns [256, 288): 2 bf16 per 32-bit column
        uint32_t v[32];
        const int row = warp * 32 + lane;
#pragma unroll
        for (int c = 0; c < 32; ++c)
            v[c] = validate ? pack_bf16((float)a_val(128 * rank + row, 2 * c), (float)a_val(128 * rank + row, 2 * c + 1)) : 0u;
        tmem_st32(tb + lane_base + 256, v);
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    unsigned long long t0 = 0, t1 = 0;
    if (warp == 0) {
        const bool leader = elect_one();
        const uint32_t idesc = idesc_bf16(256, n_dim, form == 2);
        const uint32_t a_addr = smem_u32(smem + kOffA), bk = smem_u32(smem + kOffBk), bmn = smem_u32(smem + kOffBmn);
        t0 = clock64();
        if (rank == 0) {
            for (int b = 0; b < batches; ++b) {
                if (leader) {
                    for (int i = 0; i < batch; ++i) {
                        const int kk = i & 3;
                        const uint64_t bd = form == 2 ? smem_desc_sw128(bmn + kk * 2048, kK * 128, 1024)
                                                      : smem_desc_sw128(bk + kk * 32, 16, 1024);
                        const bool acc = validate ? i > 0 : true;
                        if (form == 0) mma2_ss(tb, smem_desc_sw128(a_addr + kk * 32, 16, 1024), bd, idesc, acc);
                        else mma2_ts(tb, tb + 256 + kk * 8, bd, idesc, acc);
                    }
                    tc_commit2(&bar[b & 1]);
                }
                __syncwarp();
                if (b >= 1) mbar_wait(&bar[(b - 1) & 1], ((b - 1) >> 1) & 1);
            }
            mbar_wait(&bar[(batches - 1) & 1], ((batches - 1) >> 1) & 1);
        } else {
            for (int b = 0; b < batches; ++b) mbar_wait(&bar[b & 1], (b >> 1) & 1);  // the multicast arrivals
        }
        t1 = clock64();
        if (lane == 0) clk[blockIdx.x] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (validate) {
        int wrong = 0;
        const int row = 128 * rank + warp * 32 + lane;
        for (int c0 = 0; c0 < n_dim; c0 += 32) {
            uint32_t r[32];
            tmem_ld32(tb + lane_base + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                int want = 0;
                for (int k = 0; k < kK; ++k) want += a_val(row, k) * b_val(c0 + c, k);
                wrong += (__uint_as_float(r[c]) != (float)want);
            }
        }
        if (wrong) atomicAdd(&bad[pair], wrong);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (warp == 0) tmem_dealloc2(tb);
}

int main() {
    unsigned long long* d_clk;
    int* d_bad;
    cudaMalloc(&d_clk, 148 * 8);
    cudaMalloc(&d_bad, 74 * 4);
    cudaFuncSetAttribute(probe2, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem + 1024);
    const char* names[] = {"SS  K/K ", "TS  K   ", "TS  MN  "};
    // (1) semantics
    for (int form = 0; form < 3; ++form)
        for (int n : {128, 256}) {
            cudaMemset(d_bad, 0, 74 * 4);
            probe2<<<148, 128, kSmem + 1024>>>(form, n, 1, 4, 1, d_clk, d_bad);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("validate %s N=%d: error %s\n", names[form], n, cudaGetErrorString(e)); return 1; }
            int h[74], total = 0;
            cudaMemcpy(h, d_bad, sizeof(h), cudaMemcpyDeviceToHost);
            for (int v : h) total += v;
            printf("validate %s M=256 N=%3d K=64: %d wrong elements of %d\n", names[form], n, total, 74 * 256 * n);
        }
    // (2) rate
    for (int form = 0; form < 3; ++form)
        for (int n : {64, 128, 256})
            for (int batch : {8, 64}) {
                const int batches = 4096 / batch;
                unsigned long long h[148];
                for (int rep = 0; rep < 2; ++rep) {
                    probe2<<<148, 128, kSmem + 1024>>>(form, n, batches, batch, 0, d_clk, d_bad);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                }
                cudaMemcpy(h, d_clk, sizeof(h), cudaMemcpyDeviceToHost);
                double cyc = 0;
                for (int i = 0; i < 148; i += 2) cyc += h[i];
                cyc /= 74;
                printf("pair %s M=256 N=%3d batch=%2d: %.1f clk per MMA (per-SM share of the tile: 128 x %d x 16, ideal %d)\n",
                       names[form], n, batch, cyc / 4096, n, n / 2);
            }
    return 0;
}
