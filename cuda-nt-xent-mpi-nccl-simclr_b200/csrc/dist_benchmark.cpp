// dist_benchmark.cpp -- multi-GPU C++ benchmark of the row-partitioned NT-Xent step, on the C-ABI
// alone (no torch): one process, one host thread per GPU, ntxent_b200_comm_init_all +
// ntxent_b200_dist_forward_backward.
//
// The reference's benchmark (src/benchmark.cpp:56-90) is single-GPU and forward-only, and its
// CMake file only names NCCL/MPI (CMakeLists.txt:13-14,41-47,115-121); this is the multi-GPU half
// that north_star asks of "benchmark.cpp".  `ntxent_benchmark --gpus N ...` forwards here.
//
//   ntxent_dist_benchmark --gpus N [--M 32768] [--D 128] [--tau 0.1] [--dtype bf16|fp32]
//                         [--exchange push|nccl] [--bwd symmetric|contract] [--steps 50] [--warmup 10]
//
// Timing: every rank brackets each step with CUDA events on its stream; ranks meet at a host
// barrier before every step; a step counts as the slowest rank's time.  Prints one JSON line.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "ntxent_b200.h"

namespace {

struct Barrier {  // reusable host barrier (C++17: no std::barrier)
    explicit Barrier(int n) : n_(n) {}
    void wait() {
        std::unique_lock<std::mutex> lock(mu_);
        const unsigned gen = gen_;
        if (++count_ == n_) {
            count_ = 0;
            ++gen_;
            cv_.notify_all();
        } else {
            cv_.wait(lock, [&] { return gen != gen_; });
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    int n_, count_ = 0;
    unsigned gen_ = 0;
};

struct Options {
    int gpus = 1, steps = 50, warmup = 10;
    long long M = 32768, D = 128;
    float tau = 0.1f;
    bool bf16 = true;
    int exchange = NTXENT_EXCHANGE_PUSH, bwd = NTXENT_BWD_SYMMETRIC;
};

#define CK(expr)                                                                                  \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            fprintf(stderr, "%s:%d: %s -> %s\n", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
            failed = true;                                                                        \
        }                                                                                         \
    } while (0)

}  // namespace

int ntxent_dist_benchmark_main(int argc, char** argv) {
    Options o;
    for (int i = 1; i < argc; ++i) {
        auto arg = [&](const char* name) { return !std::strcmp(argv[i], name) && i + 1 < argc; };
        if (arg("--gpus")) o.gpus = std::atoi(argv[++i]);
        else if (arg("--steps") || arg("--runs")) o.steps = std::atoi(argv[++i]);
        else if (arg("--warmup")) o.warmup = std::atoi(argv[++i]);
        else if (arg("--M")) o.M = std::atoll(argv[++i]);
        else if (arg("--D")) o.D = std::atoll(argv[++i]);
        else if (arg("--tau")) o.tau = (float)std::atof(argv[++i]);
        else if (arg("--dtype")) o.bf16 = std::strcmp(argv[++i], "fp32") != 0;
        else if (arg("--exchange")) o.exchange = std::strcmp(argv[++i], "nccl") ? NTXENT_EXCHANGE_PUSH : NTXENT_EXCHANGE_NCCL;
        else if (arg("--bwd")) o.bwd = std::strcmp(argv[++i], "contract") ? NTXENT_BWD_SYMMETRIC : NTXENT_BWD_CONTRACT;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        fprintf(stderr, "no CUDA device: the NT-Xent engine has no CPU path\n");
        return 1;
    }
    if (o.gpus < 1 || o.gpus > ndev || o.M % o.gpus || o.steps < 1) {
        fprintf(stderr, "need 1 <= --gpus <= %d dividing M = %lld\n", ndev, o.M);
        return 1;
    }
    if (o.exchange == NTXENT_EXCHANGE_PUSH) o.bwd = NTXENT_BWD_SYMMETRIC;
    ntxent_b200_set_peer_wait_timeout_ms(20000);  // a benchmark: a rank that fell over should not cost ten minutes
    const int R = o.gpus;
    const long long rows = o.M / R;
    std::vector<int> devs(R);
    for (int r = 0; r < R; ++r) devs[r] = r;
    std::vector<ntxent_comm_t*> comms(R, nullptr);
    const int dtype = o.bf16 ? NTXENT_BF16 : NTXENT_F32;
    int rc = ntxent_b200_comm_init_all(R, devs.data(), o.M, o.D, dtype, 0, comms.data());
    if (rc != NTXENT_OK) {
        fprintf(stderr, "ntxent_b200_comm_init_all: %d (%s)\n", rc, ntxent_b200_strerror(rc));
        return 1;
    }
    // the benchmark's input: z ~ N(0,1), one generator, rank r takes its rows
    std::vector<float> z((size_t)o.M * o.D);
    {
        std::mt19937_64 gen(1234);
        std::normal_distribution<float> nd(0.f, 1.f);
        for (auto& v : z) v = nd(gen);
    }
    Barrier bar(R);
    std::vector<std::vector<float>> ms(R, std::vector<float>(o.steps, 0.f));
    std::vector<float> loss(R, 0.f);
    std::atomic<int> errors{0};
    auto work = [&](int r) {
        bool failed = false;
        CK(cudaSetDevice(devs[r]));
        const size_t n = (size_t)rows * o.D, esz = o.bf16 ? 2 : 4;
        void *dz = nullptr, *dzo = nullptr;
        float* dloss = nullptr;
        cudaStream_t st;
        CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        CK(cudaMalloc(&dz, n * esz));
        CK(cudaMalloc(&dzo, n * esz));
        CK(cudaMalloc((void**)&dloss, 256));
        const float* src = z.data() + (size_t)r * n;
        if (o.bf16) {
            std::vector<__nv_bfloat16> h(n);
            for (size_t i = 0; i < n; ++i) h[i] = __float2bfloat16(src[i]);
            CK(cudaMemcpy(dz, h.data(), n * 2, cudaMemcpyHostToDevice));
        } else {
            CK(cudaMemcpy(dz, src, n * 4, cudaMemcpyHostToDevice));
        }
        std::vector<cudaEvent_t> a(o.steps), b(o.steps);
        for (int i = 0; i < o.steps; ++i) { CK(cudaEventCreate(&a[i])); CK(cudaEventCreate(&b[i])); }
        for (int i = -o.warmup; i < o.steps; ++i) {
            CK(cudaStreamSynchronize(st));
            bar.wait();  // ranks start a step together (and a failed rank still meets the others)
            if (failed) continue;
            if (i >= 0) CK(cudaEventRecord(a[i], st));
            const int c = ntxent_b200_dist_forward_backward(comms[r], dz, nullptr, o.tau, o.exchange, o.bwd, dloss,
                                                            dzo, st);
            if (c != NTXENT_OK) {
                fprintf(stderr, "rank %d: dist_forward_backward: %d (%s)\n", r, c, ntxent_b200_strerror(c));
                failed = true;
            }
            if (i >= 0) CK(cudaEventRecord(b[i], st));
        }
        CK(cudaStreamSynchronize(st));
        if (!failed) {
            for (int i = 0; i < o.steps; ++i) CK(cudaEventElapsedTime(&ms[r][i], a[i], b[i]));
            CK(cudaMemcpy(&loss[r], dloss, 4, cudaMemcpyDeviceToHost));
            if (ntxent_b200_comm_status(comms[r]) != 0) failed = true;
        }
        if (failed) errors++;
        for (int i = 0; i < o.steps; ++i) { cudaEventDestroy(a[i]); cudaEventDestroy(b[i]); }
        cudaFree(dz); cudaFree(dzo); cudaFree(dloss);
        cudaStreamDestroy(st);
    };
    std::vector<std::thread> threads;
    for (int r = 0; r < R; ++r) threads.emplace_back(work, r);
    for (auto& t : threads) t.join();
    for (auto* c : comms) ntxent_b200_comm_destroy(c);
    if (errors) {
        fprintf(stderr, "%d rank(s) failed\n", errors.load());
        return 1;
    }
    std::vector<float> step_ms(o.steps);
    for (int i = 0; i < o.steps; ++i) {
        float m = 0.f;
        for (int r = 0; r < R; ++r) m = std::max(m, ms[r][i]);
        step_ms[i] = m;
    }
    std::sort(step_ms.begin(), step_ms.end());
    double mean = 0;
    for (float v : step_ms) mean += v;
    mean /= o.steps;
    const double flop = 8.0 * (double)o.M * (double)o.M * (double)o.D;
    bool same_loss = true;
    for (int r = 1; r < R; ++r) same_loss = same_loss && std::fabs(loss[r] - loss[0]) <= 1e-6f * std::fabs(loss[0]);
    printf("{\"benchmark\": \"ntxent_dist\", \"gpus\": %d, \"M\": %lld, \"D\": %lld, \"tau\": %g, \"dtype\": \"%s\", "
           "\"exchange\": \"%s\", \"backward\": \"%s\", \"steps\": %d, \"ms_mean\": %.5f, \"ms_median\": %.5f, "
           "\"ms_min\": %.5f, \"tflops_8M2D\": %.1f, \"loss\": %.8f, \"loss_same_on_all_ranks\": %s, "
           "\"timing\": \"per-step CUDA events, host barrier before each step, max over ranks, eager launches\"}\n",
           R, o.M, o.D, o.tau, o.bf16 ? "bf16" : "fp32", o.exchange == NTXENT_EXCHANGE_PUSH ? "push" : "nccl",
           o.bwd == NTXENT_BWD_SYMMETRIC ? "symmetric" : "contract", o.steps, mean, step_ms[o.steps / 2], step_ms[0],
           flop / (mean * 1e-3) / 1e12, loss[0], same_loss ? "true" : "false");
    return 0;
}

#ifdef NTXENT_DIST_BENCH_MAIN
int main(int argc, char** argv) { return ntxent_dist_benchmark_main(argc, argv); }
#endif
