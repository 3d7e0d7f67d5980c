// Does tcgen05.ld share TMEM read bandwidth with the tensor core's operand/accumulator traffic?
// Warps 1..N stream tcgen05.ld; thread 0 issues tcgen05.mma continuously in one of three forms.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../cuda-nt-xent-mpi-nccl-simclr_b200/csrc/tc_ptx.cuh"
using namespace ntxent_b200::ptx;

// mma_mode: 0 none, 1 SS (A,B smem), 2 TS (A tmem, B smem)
__global__ void probe(int iters, int mma_mode, int n_dim, unsigned long long* out, unsigned long long* out_mma, float* sink) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint32_t slot;
    __shared__ uint64_t bar[2];
    __shared__ volatile int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_mbar_init(); stop = 0; }
    if (warp == 0) tmem_alloc<512>(&slot);
    // zero the operand tiles so that the MMAs see finite data
    for (int i = threadIdx.x; i < 65536 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = slot;
    float acc = 0.f;
    if (warp == 0) {
        if (lane == 0 && mma_mode != 0) {
            const uint32_t idesc = idesc_bf16(128, n_dim, false);
            const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
            unsigned long long t0 = clock64();
            long long n = 0;
            int batch = 0;
            while (!stop) {
                for (int kk = 0; kk < 8; ++kk) {
                    const uint64_t bd = smem_desc_sw128(b_addr + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024);
                    if (mma_mode == 1)
                        mma_ss_f16(tb, smem_desc_sw128(a_addr + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024), bd, idesc, kk > 0);
                    else
                        mma_ts_f16(tb, tb + 256 + kk * 8, bd, idesc, kk > 0);
                }
                tc_commit(&bar[batch & 1]);
                if (batch >= 1) mbar_wait(&bar[(batch - 1) & 1], ((batch - 1) >> 1) & 1);
                ++batch; n += 8;
            }
            mbar_wait(&bar[(batch - 1) & 1], ((batch - 1) >> 1) & 1);
            unsigned long long t1 = clock64();
            out_mma[blockIdx.x * 2] = t1 - t0; out_mma[blockIdx.x * 2 + 1] = n;
        }
        __syncwarp();
    } else {
        const uint32_t base = tb + ((uint32_t)((warp & 3) * 32) << 16) + 320;
        unsigned long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            uint32_t v[32];
            tmem_ld32(base + (i & 3) * 32, v);
            tmem_ld_wait();
            acc += __uint_as_float(v[i & 31]);
        }
        unsigned long long t1 = clock64();
        if (threadIdx.x == 32) out[blockIdx.x] = t1 - t0;
        // last LDTM warp to finish stops the MMA thread (approximately: any warp finishing stops it)
        __syncwarp();
        if (lane == 0) stop = 1;
    }
    if (acc == 123.456f) sink[0] = acc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tb);
}

int main() {
    unsigned long long *d_out, *d_mma; float* d_sink;
    cudaMalloc(&d_out, 148 * 8); cudaMalloc(&d_mma, 148 * 16); cudaMalloc(&d_sink, 4);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 70 * 1024);
    const int iters = 2048;
    const char* names[] = {"no MMA", "SS MMA", "TS MMA"};
    for (int ndim : {128, 256}) {
      for (int mode = 0; mode < 3; ++mode) {
        for (int warps : {4, 16}) {
            unsigned long long h[148], hm[296];
            cudaMemset(d_mma, 0, 148 * 16);
            for (int rep = 0; rep < 2; ++rep) {
                probe<<<148, (warps + 1) * 32, 70 * 1024>>>(iters, mode, ndim, d_out, d_mma, d_sink);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            }
            cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
            cudaMemcpy(hm, d_mma, sizeof(hm), cudaMemcpyDeviceToHost);
            double cyc = 0; for (int i = 0; i < 148; ++i) cyc += h[i]; cyc /= 148;
            double mc = 0, mn = 0; for (int i = 0; i < 148; ++i) { mc += hm[2 * i]; mn += hm[2 * i + 1]; }
            printf("N=%3d %-7s ldtm warps=%2d: LDTM %.1f B/clk/SM", ndim, names[mode], warps, warps * 4096.0 * iters / cyc);
            if (mode) printf("   MMA %.1f clk per (128x%dx16) instr", mc / mn, ndim);
            printf("\n");
        }
      }
    }
    return 0;
}
