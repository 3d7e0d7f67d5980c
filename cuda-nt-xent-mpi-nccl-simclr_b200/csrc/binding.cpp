// binding.cpp -- the torch extension `ntxent_cuda`: the reference's operator surface
// (src/binding.cpp:7-10, src/binding_new.cpp:4-21) implemented over the C-ABI.
//
// One shared object is both an importable pybind module (forward / backward /
// forward_with_state / check_tensor_core_support) and a TORCH_LIBRARY, so that
// `torch.ops.ntxent_cuda.forward(z, T)` -- what python/test.py:137,71,102 calls -- resolves.
// This TU owns validation, output/workspace allocation on the caching allocator, the device
// guard and the current stream; all compute is behind ntxent_b200_* calls.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include "ntxent_b200.h"
#include "ntxent_kernel.cuh"

namespace ntxent {
namespace {

void check_rc(int rc, const char* what) {
    TORCH_CHECK(rc == NTXENT_OK, what, " failed: ", ntxent_b200_strerror(rc), " (code ", rc, ")");
}

struct Prepared {
    torch::Tensor z, workspace;
    int64_t M, D;
    int dtype;
    unsigned flags;
};

Prepared prepare(const torch::Tensor& z_in, double T, bool use_mixed_precision) {
    TORCH_CHECK(z_in.is_cuda(), "ntxent: z must be a CUDA tensor (there is no CPU path)");
    TORCH_CHECK(z_in.dim() == 2, "ntxent: z must be [2N, D], got ", z_in.dim(), " dimensions");
    TORCH_CHECK(z_in.scalar_type() == at::kFloat || z_in.scalar_type() == at::kBFloat16,
                "ntxent: z must be float32 or bfloat16, got ", z_in.scalar_type());
    TORCH_CHECK(T > 0.0, "ntxent: temperature must be positive, got ", T);
    Prepared p;
    p.z = z_in.contiguous();
    p.M = p.z.size(0);
    p.D = p.z.size(1);
    TORCH_CHECK(p.M >= 2 && p.M % 2 == 0, "ntxent: z needs an even number of rows (two stacked views), got ", p.M);
    p.dtype = p.z.scalar_type() == at::kBFloat16 ? NTXENT_BF16 : NTXENT_F32;
    p.flags = use_mixed_precision ? NTXENT_FLAG_MIXED_PRECISION : 0u;
    const size_t bytes = ntxent_b200_workspace_bytes(p.M, p.D, p.dtype, p.flags);
    TORCH_CHECK(bytes > 0, "ntxent: unsupported shape [", p.M, ", ", p.D,
                "] (need D % 8 == 0 and 8 <= D <= 1024)");
    p.workspace = torch::empty({static_cast<int64_t>(bytes)}, p.z.options().dtype(at::kByte));
    return p;
}

std::tuple<torch::Tensor, torch::Tensor> forward_impl(const torch::Tensor& z_in, double T, bool mp) {
    TORCH_CHECK(z_in.is_cuda(), "ntxent: z must be a CUDA tensor (there is no CPU path)");
    c10::cuda::CUDAGuard guard(z_in.device());
    Prepared p = prepare(z_in, T, mp);
    auto fopt = p.z.options().dtype(at::kFloat);
    torch::Tensor loss = torch::empty({1}, fopt), lse = torch::empty({p.M}, fopt);
    check_rc(ntxent_b200_forward(p.z.data_ptr(), p.M, p.D, static_cast<float>(T), p.dtype, p.flags,
                                 p.workspace.data_ptr(), p.workspace.numel(), loss.data_ptr<float>(),
                                 lse.data_ptr<float>(), at::cuda::getCurrentCUDAStream()),
             "ntxent_b200_forward");
    return {loss, lse};
}

std::tuple<torch::Tensor, torch::Tensor> backward_impl(const torch::Tensor& z_in,
                                                       const torch::Tensor& lse_in,
                                                       const torch::Tensor& grad_out, double T,
                                                       bool mp) {
    TORCH_CHECK(z_in.is_cuda(), "ntxent: z must be a CUDA tensor (there is no CPU path)");
    c10::cuda::CUDAGuard guard(z_in.device());
    Prepared p = prepare(z_in, T, mp);
    TORCH_CHECK(lse_in.is_cuda() && lse_in.numel() == p.M,
                "ntxent backward: the second argument is the forward's row log-sum-exp with ", p.M,
                " entries (from ntxent_forward_with_state_cuda); got a tensor of ", lse_in.numel(),
                " elements -- an [M, M] softmax is never formed");
    torch::Tensor lse = lse_in.reshape({p.M}).to(at::kFloat).contiguous();
    torch::Tensor g;
    const float* gptr = nullptr;
    if (grad_out.defined() && grad_out.numel() > 0) {
        TORCH_CHECK(grad_out.is_cuda(), "ntxent backward: grad_out must be a CUDA tensor");
        g = grad_out.reshape({-1}).slice(0, 0, 1).to(at::kFloat).contiguous();
        gptr = g.data_ptr<float>();
    }
    torch::Tensor dz = torch::empty_like(p.z);
    torch::Tensor pos_prob = torch::empty({p.M}, p.z.options().dtype(at::kFloat));
    check_rc(ntxent_b200_backward(p.z.data_ptr(), lse.data_ptr<float>(), gptr, p.M, p.D,
                                  static_cast<float>(T), p.dtype, p.flags, p.workspace.data_ptr(),
                                  p.workspace.numel(), dz.data_ptr(), pos_prob.data_ptr<float>(),
                                  at::cuda::getCurrentCUDAStream()),
             "ntxent_b200_backward");
    return {dz, pos_prob};
}

class NTXentFunction : public torch::autograd::Function<NTXentFunction> {
public:
    static torch::Tensor forward(torch::autograd::AutogradContext* ctx, torch::Tensor z, double T,
                                 bool mp) {
        at::AutoDispatchBelowADInplaceOrView no_grad_below;
        auto [loss, lse] = forward_impl(z, T, mp);
        ctx->save_for_backward({z, lse});
        ctx->saved_data["T"] = T;
        ctx->saved_data["mp"] = mp;
        return loss;
    }
    static torch::autograd::tensor_list backward(torch::autograd::AutogradContext* ctx,
                                                 torch::autograd::tensor_list grad_outputs) {
        auto saved = ctx->get_saved_variables();
        auto [dz, pos_prob] = backward_impl(saved[0], saved[1], grad_outputs[0],
                                            ctx->saved_data["T"].toDouble(),
                                            ctx->saved_data["mp"].toBool());
        (void)pos_prob;
        return {dz, torch::Tensor(), torch::Tensor()};
    }
};

}  // namespace

torch::Tensor ntxent_forward_cuda(torch::Tensor z, double T, bool use_mixed_precision) {
    if (z.requires_grad() && at::GradMode::is_enabled())
        return NTXentFunction::apply(z, T, use_mixed_precision);
    return std::get<0>(forward_impl(z, T, use_mixed_precision));
}

std::tuple<torch::Tensor, torch::Tensor> ntxent_forward_with_state_cuda(torch::Tensor z, double T,
                                                                        bool use_mixed_precision) {
    return forward_impl(z.detach(), T, use_mixed_precision);
}

std::tuple<torch::Tensor, torch::Tensor> ntxent_backward_cuda(torch::Tensor z, torch::Tensor softmax,
                                                              torch::Tensor grad_out, double T,
                                                              bool use_mixed_precision) {
    return backward_impl(z.detach(), softmax.detach(), grad_out.detach(), T, use_mixed_precision);
}

namespace utils {
bool check_tensor_core_support() { return ntxent_b200_device_supported() == 1; }
}  // namespace utils

}  // namespace ntxent

TORCH_LIBRARY(ntxent_cuda, m) {
    m.def("forward(Tensor z, float T, bool use_mixed_precision=False) -> Tensor",
          &ntxent::ntxent_forward_cuda);
    m.def("forward_with_state(Tensor z, float T, bool use_mixed_precision=False) -> (Tensor, Tensor)",
          &ntxent::ntxent_forward_with_state_cuda);
    m.def("backward(Tensor z, Tensor softmax, Tensor grad_out, float T, bool "
          "use_mixed_precision=False) -> (Tensor, Tensor)",
          &ntxent::ntxent_backward_cuda);
    m.def("check_tensor_core_support() -> bool", &ntxent::utils::check_tensor_core_support);
}

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
    m.def("forward", &ntxent::ntxent_forward_cuda, "NT-Xent forward (CUDA)", py::arg("z"), py::arg("T"),
          py::arg("use_mixed_precision") = false);
    m.def("forward_with_state", &ntxent::ntxent_forward_with_state_cuda,
          "NT-Xent forward returning (loss, lse)", py::arg("z"), py::arg("T"),
          py::arg("use_mixed_precision") = false);
    m.def("backward", &ntxent::ntxent_backward_cuda, "NT-Xent backward (CUDA)", py::arg("z"),
          py::arg("softmax"), py::arg("grad_out"), py::arg("T"), py::arg("use_mixed_precision") = false);
    m.def("check_tensor_core_support", &ntxent::utils::check_tensor_core_support,
          "Check whether the current GPU runs the tcgen05 kernels");
}
