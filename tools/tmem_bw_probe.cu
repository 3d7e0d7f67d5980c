// Microbenchmark: tcgen05.ld (TMEM -> registers) and MUFU.EX2 / FFMA issue throughput per SM.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bw_probe tmem_bw_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../cuda-nt-xent-mpi-nccl-simclr_b200/csrc/tc_ptx.cuh"
using namespace ntxent_b200::ptx;

__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&r)[64]) {
    uint32_t (&a)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[0]);
    uint32_t (&b)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[32]);
    tmem_ld32(taddr, a);
    tmem_ld32(taddr + 32, b);
}

template <int MODE>
__global__ void probe(int iters, unsigned long long* out, float* sink, float k) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc<512>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    float acc = 0.f;
    __syncthreads();
    const unsigned long long t0 = clock64();
    if (MODE == 0) {  // LDTM x32 back to back, one wait per load
        for (int i = 0; i < iters; ++i) {
            uint32_t v[32];
            tmem_ld32(base + (i & 7) * 32, v);
            tmem_ld_wait();
            uint32_t x = 0;
#pragma unroll
            for (int e = 0; e < 32; ++e) x ^= v[e];
            acc += __uint_as_float(x);
        }
    } else if (MODE == 1) {  // two x32 loads per wait
        for (int i = 0; i < iters; i += 2) {
            uint32_t v[64];
            tmem_ld64(base + (i & 6) * 32, v);
            tmem_ld_wait();
            uint32_t x = 0;
#pragma unroll
            for (int e = 0; e < 64; ++e) x ^= v[e];
            acc += __uint_as_float(x);
        }
    } else if (MODE == 2) {  // MUFU.EX2 throughput: 32 independent per iteration
        float x[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) x[e] = k * (float)(e + lane);
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int e = 0; e < 32; ++e) x[e] = ex2(x[e]) - 1.5f;
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) acc += x[e];
    } else if (MODE == 3) {  // FFMA throughput (3-register form)
        float x[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) x[e] = k * (float)(e + lane);
        float a = k, b = k * 0.5f;
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int e = 0; e < 32; ++e) x[e] = fmaf(x[e], a, b);
        }
#pragma unroll
        for (int e = 0; e < 32; ++e) acc += x[e];
    } else if (MODE == 4) {  // the forward mix: LDTM + FFMA + EX2 + FADD
        for (int i = 0; i < iters; ++i) {
            uint32_t v[32];
            tmem_ld32(base + (i & 7) * 32, v);
            tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 32; ++e) acc += ex2(fmaf(__uint_as_float(v[e]), k, -k));
        }
    } else if (MODE == 5) {  // STTM x16
        uint32_t p[16];
#pragma unroll
        for (int e = 0; e < 16; ++e) p[e] = lane + e;
        for (int i = 0; i < iters; ++i) {
            tmem_st16(base + (i & 15) * 16, p);
            tmem_st_wait();
        }
    }
    const unsigned long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 123.456f) sink[0] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(slot);
}

int main() {
    unsigned long long* d_out; float* d_sink;
    cudaMalloc(&d_out, 148 * 8); cudaMalloc(&d_sink, 4);
    const int iters = 4096;
    const char* names[] = {"LDTM x32 (1 per wait)", "LDTM x32 (2 per wait)", "MUFU.EX2", "FFMA", "fwd mix", "STTM x16"};
    for (int mode = 0; mode < 6; ++mode) {
        for (int warps : {4, 8, 16}) {
            unsigned long long h[148];
            for (int rep = 0; rep < 2; ++rep) {
                switch (mode) {
                    case 0: probe<0><<<148, warps * 32>>>(iters, d_out, d_sink, 0.001f); break;
                    case 1: probe<1><<<148, warps * 32>>>(iters, d_out, d_sink, 0.001f); break;
                    case 2: probe<2><<<148, warps * 32>>>(iters, d_out, d_sink, 0.001f); break;
                    case 3: probe<3><<<148, warps * 32>>>(iters, d_out, d_sink, 0.001f); break;
                    case 4: probe<4><<<148, warps * 32>>>(iters, d_out, d_sink, 0.001f); break;
                    case 5: probe<5><<<148, warps * 32>>>(iters, d_out, d_sink, 0.001f); break;
                }
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
            double cyc = 0; for (int i = 0; i < 148; ++i) cyc += h[i]; cyc /= 148;
            double per_iter = cyc / iters;
            if (mode <= 1 || mode == 4)
                printf("%-24s warps=%2d  %.1f clk/iter/warp  -> %.1f B/clk/SM  (%.2f elem/clk/SM)\n", names[mode], warps,
                       per_iter, warps * 32 * 32 * 4 / per_iter, warps * 32 * 32 / per_iter);
            else if (mode == 5)
                printf("%-24s warps=%2d  %.1f clk/iter/warp  -> %.1f B/clk/SM\n", names[mode], warps, per_iter,
                       warps * 32 * 16 * 4 / per_iter);
            else
                printf("%-24s warps=%2d  %.1f clk/iter/warp  -> %.2f ops/clk/SM\n", names[mode], warps, per_iter,
                       warps * 32 * 32 / per_iter);
        }
    }
    return 0;
}
