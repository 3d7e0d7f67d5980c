// ntxent_kernel.cuh -- C++ operator surface of the B200-native NT-Xent engine.
//
// Same file name, namespace and function names as the reference's header
// (/root/reference/include/ntxent_kernel.cuh:24-52) so that its call sites -- src/binding.cpp:4-9,
// src/benchmark.cpp:26,32, tests/test_forward.cpp:23,34,47, tests/test_backward.cpp:24-28 --
// compile against this library.  Nothing else is shared: there are no kernel declarations,
// no cuBLAS handle and no launch constants here; the kernels sit behind the C-ABI in
// ntxent_b200.h and this header is plain C++ (a host compiler is enough).
//
// Semantics that differ from the reference are listed in INTEGRATION.md; in short z is the
// stacked [2N, D] batch, backward takes the forward's row log-sum-exp instead of an [M, M]
// softmax, and grad_out is honoured.
#pragma once
#include <torch/extension.h>

#include <tuple>

namespace ntxent {

// NT-Xent forward.  z: CUDA [M, D], M even, float32 or bfloat16.  Returns the mean loss as a
// [1] float32 tensor; differentiable when z requires grad (tests/test_forward.cpp:29-38).
// use_mixed_precision: bf16 operands with fp32 accumulation even for float32 z.
torch::Tensor ntxent_forward_cuda(torch::Tensor z, double T, bool use_mixed_precision = false);

// Forward that also returns the state backward needs: (loss[1], lse[M]).  This is what the
// reference's tests mean by std::get<0>(outputs) (tests/test_backward.cpp:24-25), which cannot
// compile against the declared single-tensor return type.
std::tuple<torch::Tensor, torch::Tensor> ntxent_forward_with_state_cuda(
    torch::Tensor z, double T, bool use_mixed_precision = false);

// NT-Xent backward.  `softmax` is the forward's saved state lse[M] (an [M, M] tensor is
// rejected); grad_out: the incoming gradient of the loss (first element is used).
// Returns (grad_z[M, D] in z's dtype, grad_internal[M] = P_{i,pos(i)}).
std::tuple<torch::Tensor, torch::Tensor> ntxent_backward_cuda(torch::Tensor z, torch::Tensor softmax,
                                                              torch::Tensor grad_out, double T,
                                                              bool use_mixed_precision = false);

namespace utils {
// True when the current device runs the tcgen05 kernels (compute capability 10.x).
// Reference: include/ntxent_kernel.cuh:99-110 (cc.major >= 7).
bool check_tensor_core_support();
}  // namespace utils

}  // namespace ntxent

// The reference's tests and benchmark call these unqualified although they are declared inside
// namespace ntxent; exporting them keeps those call sites compiling.
using ntxent::ntxent_backward_cuda;
using ntxent::ntxent_forward_cuda;
using ntxent::ntxent_forward_with_state_cuda;
