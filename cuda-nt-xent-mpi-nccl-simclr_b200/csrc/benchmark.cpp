// benchmark.cpp -- C++ benchmark of the NT-Xent operator (the reference's src/benchmark.cpp role).
//
// The reference times the forward only, with a host clock around a device synchronise, over
// rows in {32..1024} x D in {64,128,256} at T = 0.07 (src/benchmark.cpp:30-39,68-71) and prints
// mean/std/min/max.  This one keeps that sweep and table, and adds what BASELINE.json asks for:
// forward+backward, CUDA-event timing, TFLOP/s on the 8 M^2 D work model, the BASELINE configs,
// bf16 inputs and a JSON report.
//
//   ntxent_benchmark [--runs N] [--json FILE] [--quick]
//   ntxent_benchmark --gpus N [--M ..] [--D ..] [--exchange push|nccl] [--bwd symmetric|contract] ...
//       the row-partitioned step on N GPUs of this box through the C-ABI communicator
//       (dist_benchmark.cpp; one process, one host thread per GPU)
#include <ATen/cuda/CUDAContext.h>
#include <cuda_runtime.h>
#include <torch/torch.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <numeric>
#include <sstream>
#include <vector>

#include "ntxent_kernel.cuh"

namespace {

struct Stats {
    double mean = 0, stddev = 0, min = 0, max = 0;
};

Stats summarise(const std::vector<float>& ms) {
    Stats s;
    s.mean = std::accumulate(ms.begin(), ms.end(), 0.0) / ms.size();
    double sq = 0;
    for (float v : ms) sq += (v - s.mean) * (v - s.mean);
    s.stddev = std::sqrt(sq / ms.size());
    s.min = *std::min_element(ms.begin(), ms.end());
    s.max = *std::max_element(ms.begin(), ms.end());
    return s;
}

struct Row {
    int64_t rows, dim;
    double T;
    std::string dtype;
    Stats fwd, fwd_bwd;
    double tflops;
};

template <class Fn>
std::vector<float> time_ms(Fn&& fn, int warmup, int runs) {
    auto stream = at::cuda::getCurrentCUDAStream();
    for (int i = 0; i < warmup; ++i) fn();
    std::vector<cudaEvent_t> a(runs), b(runs);
    for (int i = 0; i < runs; ++i) { cudaEventCreate(&a[i]); cudaEventCreate(&b[i]); }
    cudaStreamSynchronize(stream);
    for (int i = 0; i < runs; ++i) {
        cudaEventRecord(a[i], stream);
        fn();
        cudaEventRecord(b[i], stream);
    }
    cudaStreamSynchronize(stream);
    std::vector<float> ms(runs);
    for (int i = 0; i < runs; ++i) {
        cudaEventElapsedTime(&ms[i], a[i], b[i]);
        cudaEventDestroy(a[i]);
        cudaEventDestroy(b[i]);
    }
    return ms;
}

Row measure(int64_t rows, int64_t dim, double T, bool bf16, int runs) {
    torch::manual_seed(1234);
    auto z = torch::randn({rows, dim}, torch::TensorOptions().device(torch::kCUDA));
    if (bf16) z = z.to(torch::kBFloat16);
    auto one = torch::ones({1}, torch::TensorOptions().device(torch::kCUDA));
    Row r{rows, dim, T, bf16 ? "bf16" : "fp32", {}, {}, 0};
    r.fwd = summarise(time_ms([&] { ntxent_forward_cuda(z, T); }, 3, runs));
    r.fwd_bwd = summarise(time_ms(
        [&] {
            auto [loss, state] = ntxent_forward_with_state_cuda(z, T);
            ntxent_backward_cuda(z, state, one, T);
        },
        3, runs));
    r.tflops = 8.0 * rows * rows * dim / (r.fwd_bwd.mean * 1e-3) / 1e12;
    return r;
}

void print_row(const Row& r) {
    std::cout << std::setw(8) << r.rows << std::setw(6) << r.dim << std::setw(6) << r.dtype << std::fixed
              << std::setprecision(4) << std::setw(11) << r.fwd.mean << std::setw(10) << r.fwd.stddev
              << std::setw(10) << r.fwd.min << std::setw(10) << r.fwd.max << std::setw(12) << r.fwd_bwd.mean
              << std::setprecision(1) << std::setw(11) << r.tflops << std::endl;
}

}  // namespace

int ntxent_dist_benchmark_main(int argc, char** argv);  // dist_benchmark.cpp

int main(int argc, char** argv) {
    for (int i = 1; i < argc; ++i)
        if (!std::strcmp(argv[i], "--gpus")) return ntxent_dist_benchmark_main(argc, argv);
    int runs = 100;
    bool quick = false;
    std::string json_path;
    for (int i = 1; i < argc; ++i) {
        if (!std::strcmp(argv[i], "--runs") && i + 1 < argc) runs = std::atoi(argv[++i]);
        else if (!std::strcmp(argv[i], "--json") && i + 1 < argc) json_path = argv[++i];
        else if (!std::strcmp(argv[i], "--quick")) quick = true;
    }
    if (!torch::cuda::is_available()) {
        std::cerr << "CUDA is not available: the NT-Xent engine has no CPU path" << std::endl;
        return 1;
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    std::cout << "NT-Xent operator benchmark on " << prop.name << " (sm_" << prop.major << prop.minor << ", "
              << prop.multiProcessorCount << " SMs), tcgen05 kernels "
              << (ntxent::utils::check_tensor_core_support() ? "available" : "NOT available") << "\n"
              << "times in ms over " << runs << " runs (CUDA events); TFLOP/s = 8 M^2 D / (fwd+bwd mean)\n"
              << std::setw(8) << "rows" << std::setw(6) << "D" << std::setw(6) << "type" << std::setw(11)
              << "fwd mean" << std::setw(10) << "std" << std::setw(10) << "min" << std::setw(10) << "max"
              << std::setw(12) << "fwd+bwd" << std::setw(11) << "TFLOP/s" << std::endl;
    std::vector<Row> out;
    // the reference's sweep (src/benchmark.cpp:68-71), fp32 inputs, T = 0.07
    std::vector<int64_t> ref_rows = {32, 64, 128, 256, 512, 1024}, ref_dims = {64, 128, 256};
    if (quick) { ref_rows = {32, 1024}; ref_dims = {128}; }
    for (int64_t r : ref_rows)
        for (int64_t d : ref_dims) {
            out.push_back(measure(r, d, 0.07, false, runs));
            print_row(out.back());
        }
    // BASELINE.json configurations
    struct Cfg { int64_t rows, dim; double T; bool bf16; };
    std::vector<Cfg> cfgs = {{512, 128, 0.5, false}, {4096, 128, 0.1, true}, {8192, 128, 0.5, false},
                             {32768, 128, 0.1, true}};
    if (!quick) cfgs.push_back({131072, 256, 0.1, true});
    for (auto& c : cfgs) {
        out.push_back(measure(c.rows, c.dim, c.T, c.bf16, std::max(5, runs / 10)));
        print_row(out.back());
    }
    if (!json_path.empty()) {
        std::ofstream f(json_path);
        f << "[\n";
        for (size_t i = 0; i < out.size(); ++i) {
            const Row& r = out[i];
            f << "  {\"rows\": " << r.rows << ", \"dim\": " << r.dim << ", \"T\": " << r.T << ", \"dtype\": \""
              << r.dtype << "\", \"fwd_ms_mean\": " << r.fwd.mean << ", \"fwd_ms_std\": " << r.fwd.stddev
              << ", \"fwd_ms_min\": " << r.fwd.min << ", \"fwd_ms_max\": " << r.fwd.max
              << ", \"fwd_bwd_ms_mean\": " << r.fwd_bwd.mean << ", \"tflops\": " << r.tflops << "}"
              << (i + 1 < out.size() ? "," : "") << "\n";
        }
        f << "]\n";
    }
    return 0;
}
