// How fast does one elected lane issue/execute tcgen05.mma (M=128, K=16) in the forms the tile
// kernels use?  Zero-filled operands; one CTA per SM; a commit + wait every `batch` MMAs.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../cuda-nt-xent-mpi-nccl-simclr_b200/csrc/tc_ptx.cuh"
using namespace ntxent_b200::ptx;

// form: 0 SS K-major/K-major, 1 TS (A tmem) B K-major, 2 TS B MN-major
__global__ void probe(int form, int n_dim, int batches, int batch, unsigned long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint32_t slot;
    __shared__ uint64_t bar[2];
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc<512>(&slot);
    for (int i = threadIdx.x; i < 131072 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = slot;
    if (warp == 0) {
        const bool leader = elect_one();
        const uint32_t idesc = idesc_bf16(128, n_dim, form == 2);
        const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
        const unsigned long long t0 = clock64();
        for (int b = 0; b < batches; ++b) {
            if (leader) {
                for (int i = 0; i < batch; ++i) {
                    const int kk = i & 7;
                    const uint64_t bd = form == 2 ? smem_desc_sw128(b_addr + kk * 2048, 16384, 1024)
                                                  : smem_desc_sw128(b_addr + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024);
                    if (form == 0)
                        mma_ss_f16(tb, smem_desc_sw128(a_addr + (kk >> 2) * 16384 + (kk & 3) * 32, 16, 1024), bd, idesc, i > 0);
                    else
                        mma_ts_f16(tb, tb + 256 + kk * 8, bd, idesc, i > 0);
                }
                tc_commit(&bar[b & 1]);
            }
            __syncwarp();
            if (b >= 1) mbar_wait(&bar[(b - 1) & 1], ((b - 1) >> 1) & 1);
        }
        mbar_wait(&bar[(batches - 1) & 1], ((batches - 1) >> 1) & 1);
        const unsigned long long t1 = clock64();
        if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tb);
}

int main() {
    unsigned long long* d_out;
    cudaMalloc(&d_out, 148 * 8);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 140 * 1024);
    const char* names[] = {"SS  K/K ", "TS  K   ", "TS  MN  "};
    for (int form = 0; form < 3; ++form)
        for (int n : {64, 128})
            for (int batch : {8, 64}) {
                const int batches = 4096 / batch;
                unsigned long long h[148];
                for (int rep = 0; rep < 2; ++rep) {
                    probe<<<148, 128, 140 * 1024>>>(form, n, batches, batch, d_out);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                }
                cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
                double cyc = 0; for (int i = 0; i < 148; ++i) cyc += h[i]; cyc /= 148;
                printf("%s N=%3d batch=%2d: %.1f clk per MMA (ideal %d)\n", names[form], n, batch, cyc / 4096, n / 2);
            }
    return 0;
}
