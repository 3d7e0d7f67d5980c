"""The reference's operator surface, in Python, over the C-ABI.

Mirrors `forward(z, T, use_mixed_precision=False)` / `backward(z, softmax, grad_out, T,
use_mixed_precision=False)` / `check_tensor_core_support()` of the reference's pybind module
(src/binding.cpp:7-10, src/binding_new.cpp:4-21; C++ signatures include/ntxent_kernel.cuh:31-52).
The C++ twin of this file is csrc/binding.cpp.

Differences a caller of the reference has to know:
  * z is the stacked [2N, D] batch (the reference duplicates z[B, D] itself,
    src/ntxent_kernel.cu:161, which is not NT-Xent);
  * backward's second argument is the forward's saved state: the row log-sum-exp l[M] from
    `forward_with_state`, never an [M, M] softmax;
  * `grad_out` is honoured (the reference ignores it, src/ntxent_kernel.cu:205-239);
  * the second return value of backward is P_{i,pos(i)} as an [M] tensor.
"""
from __future__ import annotations

import torch

from . import capi


def check_tensor_core_support() -> bool:
    """True when the current device can run the tcgen05 kernels (sm_100).

    Reference: ntxent::utils::check_tensor_core_support (include/ntxent_kernel.cuh:99-110)."""
    return bool(capi.lib().ntxent_b200_device_supported())


def forward_with_state(z: torch.Tensor, T: float, use_mixed_precision: bool = False):
    """-> (loss[1], lse[M]).  `lse` is what backward needs."""
    return capi.forward(z, T, use_mixed_precision)


def forward(z: torch.Tensor, T: float, use_mixed_precision: bool = False) -> torch.Tensor:
    """NT-Xent forward: loss as a [1] fp32 tensor.  Differentiable when z requires grad
    (tests/test_forward.cpp:29-38 calls loss.backward())."""
    if z.requires_grad and torch.is_grad_enabled():
        return _NTXentFunction.apply(z, float(T), bool(use_mixed_precision))
    return capi.forward(z, T, use_mixed_precision)[0]


def backward(z: torch.Tensor, softmax: torch.Tensor, grad_out: torch.Tensor, T: float,
             use_mixed_precision: bool = False):
    """NT-Xent backward -> (grad_z[M,D], grad_internal[M]).

    `softmax` is the forward's state l[M] (see module docstring); `grad_out` may be a scalar
    tensor or any tensor whose first element is the incoming gradient of the loss."""
    return capi.backward(z, softmax, grad_out, T, use_mixed_precision)


class _NTXentFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, T, use_mixed_precision, flags=0):
        loss, lse = capi.forward(z.detach(), T, use_mixed_precision, flags)
        ctx.save_for_backward(z.detach(), lse)
        ctx.T, ctx.mp, ctx.flags = T, use_mixed_precision, flags
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        z, lse = ctx.saved_tensors
        dz, _ = capi.backward(z, lse, grad_loss, ctx.T, ctx.mp, ctx.flags, want_pos_prob=False)
        return dz, None, None, None


class _NTXentLearnableTauFunction(torch.autograd.Function):
    """Same op with the temperature as a differentiable 0-dim tensor.  The
    kernels take tau as a host scalar, so forward reads it back once (one host sync per step)."""

    @staticmethod
    def forward(ctx, z, tau, use_mixed_precision):
        T = float(tau.detach())
        if not T > 0.0:
            raise capi.NTXentError(f"temperature must stay positive, got {T}")
        loss, lse = capi.forward(z.detach(), T, use_mixed_precision)
        ctx.save_for_backward(z.detach(), lse)
        ctx.T, ctx.mp, ctx.tau_meta = T, use_mixed_precision, (tau.dtype, tau.device, tau.shape)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        z, lse = ctx.saved_tensors
        dz, rows = capi.backward_tau(z, lse, grad_loss, ctx.T, ctx.mp)
        dtype, device, shape = ctx.tau_meta
        dtau = (grad_loss.reshape(-1)[:1].float() * rows.double().sum().float()).reshape(shape)
        return dz, dtau.to(device=device, dtype=dtype), None


class _NTXentPositivesFunction(torch.autograd.Function):
    """NT-Xent with the positive of every row given by an index vector.
    Forward and backward run as one fused call in forward(); backward scales the stored gradient."""

    @staticmethod
    def forward(ctx, z, pos_lists, T, use_mixed_precision):
        need = z.requires_grad
        loss, _, dz = capi.forward_backward_pos(z.detach(), pos_lists, T, None, use_mixed_precision, want_dz=need)
        ctx.dz = dz
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        dz = ctx.dz
        return (dz.float() * grad_loss.reshape(-1)[:1].float()).to(dz.dtype), None, None, None


class GraphedNTXent:
    """Fused forward+backward for a fixed [M, D] shape, captured once into a CUDA graph and replayed
    (SURVEY.md section 8f-2): for projection-head sized batches (2N ~ 4096) the six launches of a
    step cost more host time than the kernels take.  `__call__(z)` copies z into the static input
    and returns the graph's static (loss[1], dz[M, D]); consume them before the next call."""

    def __init__(self, M: int, D: int, temperature: float, dtype: torch.dtype = torch.bfloat16,
                 device="cuda", use_mixed_precision: bool = False, flags: int = 0):
        self.T, self.mp = float(temperature), bool(use_mixed_precision)
        self.z = torch.zeros(M, D, dtype=dtype, device=device)
        self.ws = capi.workspace(M, D, self.z, flags | (capi.FLAG_MIXED_PRECISION if self.mp else 0))
        self.out = (torch.empty(1, dtype=torch.float32, device=device),
                    torch.empty(M, dtype=torch.float32, device=device), torch.empty_like(self.z))
        side = torch.cuda.Stream(device=device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                capi.forward_backward(self.z, self.T, None, self.mp, flags, self.ws, self.out)
        torch.cuda.current_stream().wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            capi.forward_backward(self.z, self.T, None, self.mp, flags, self.ws, self.out)

    def __call__(self, z: torch.Tensor):
        self.z.copy_(z, non_blocking=True)
        self.graph.replay()
        return self.out[0], self.out[2]


def ntxent_loss(z: torch.Tensor, temperature: float = 0.07,
                use_mixed_precision: bool = False, positives: torch.Tensor | None = None) -> torch.Tensor:
    """Scalar (0-dim) differentiable NT-Xent loss of the stacked batch z[2N, D].  A 0-dim tensor
    temperature that requires grad receives d loss / d tau.  `positives`: an int tensor [M] naming
    the positive of every row (default: row i's positive is row (i + N) mod 2N); pass the result
    of capi.inverse_positives() instead to skip the per-call sort."""
    if positives is not None:
        lists = positives if isinstance(positives, tuple) else capi.inverse_positives(positives.to(z.device))
        return _NTXentPositivesFunction.apply(z, lists, float(temperature), use_mixed_precision).squeeze(0)
    if isinstance(temperature, torch.Tensor):
        if temperature.requires_grad:
            return _NTXentLearnableTauFunction.apply(z, temperature, use_mixed_precision).squeeze(0)
        temperature = float(temperature)
    return forward(z, temperature, use_mixed_precision).squeeze(0)


def two_tower_loss(q: torch.Tensor, k: torch.Tensor, temperature: float = 0.07,
                   use_mixed_precision: bool = False) -> torch.Tensor:
    """Symmetric InfoNCE between two towers (CLIP's loss; SURVEY.md section 8 f4, "Q != K"):
    (CE(q k^T / T, diag) + CE(k q^T / T, diag)) / 2 over L2-normalised rows, q[i] and k[i] being
    the positive pair.  It is NT-Xent over the stacked [q; k] with the denominators restricted to
    the other tower (NTXENT_FLAG_CROSS_VIEW), so it runs on the same kernels, visiting only the
    cross-tower tiles.  0-dim, differentiable in q and k."""
    if q.shape != k.shape or q.dim() != 2:
        raise capi.NTXentError(f"two_tower_loss expects two [N, D] tensors of one shape, got {tuple(q.shape)} "
                               f"and {tuple(k.shape)}")
    z = torch.cat([q, k], dim=0)
    if z.requires_grad and torch.is_grad_enabled():
        return _NTXentFunction.apply(z, float(temperature), bool(use_mixed_precision), capi.FLAG_CROSS_VIEW).squeeze(0)
    return capi.forward(z, float(temperature), use_mixed_precision, capi.FLAG_CROSS_VIEW)[0].squeeze(0)


class NTXentLoss(torch.nn.Module):
    """nn.Module face of the op.  With `process_group` set, the batch is
    row-partitioned across the group (see distributed.py)."""

    def __init__(self, temperature: float = 0.07, use_mixed_precision: bool = False,
                 process_group=None, backward_mode: str = "symmetric",
                 learnable_temperature: bool = False):
        super().__init__()
        if learnable_temperature:
            if process_group is not None:
                raise ValueError("learnable_temperature is single-GPU only for now")
            self.temperature = torch.nn.Parameter(torch.tensor(float(temperature)))
        else:
            self.temperature = float(temperature)
        self.use_mixed_precision = bool(use_mixed_precision)
        self.process_group = process_group
        self.backward_mode = backward_mode

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        if self.process_group is not None:
            from .distributed import distributed_ntxent_loss
            return distributed_ntxent_loss(z, self.temperature, self.use_mixed_precision,
                                           group=self.process_group, backward_mode=self.backward_mode)
        if isinstance(self.temperature, torch.Tensor):
            return _NTXentLearnableTauFunction.apply(z, self.temperature, self.use_mixed_precision).squeeze(0)
        return ntxent_loss(z, self.temperature, self.use_mixed_precision)
