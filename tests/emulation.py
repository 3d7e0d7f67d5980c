"""A plain PyTorch fp32 restatement of what the bf16 tcgen05 kernels compute, ROUNDING STEP BY
ROUNDING STEP (test infrastructure, GPU only).

The CPU oracle answers "is this NT-Xent?" -- but at bf16 tolerance (1e-2 of the largest gradient
entry) it cannot see a dropped or double-counted 128 x 128 tile of the 2N = 32768 problem, whose
contribution to dz is ~3e-4.  This restatement rounds where the kernels round (U to bf16 by the
library's own normalise kernel, k*U_a to bf16 again, the weights W to bf16 before the second
contraction) and is exact fp32 everywhere else, so that what is left between it and the kernels
is accumulation order, the ex2 approximations and nothing else: ~1e-5.  Row blocks of
`block` rows keep the logits of one block (block x M fp32) in memory at a time.
"""
import math

import torch

from ntxent_b200 import capi


@torch.no_grad()
def emulate_bf16_step(z_t: torch.Tensor, tau: float, grad_out: float = 1.0, block: int = 4096,
                      sym_forward=None):
    """-> (loss float, lse[M] fp64 cpu-able tensor, dz[M, D] fp32) for bf16-operand configurations
    in the fixed-max regime (log2(e)/tau <= 50).

    sym_forward: the whole-batch forward of M >= 4096 rows runs over the upper triangle of tiles
    (csrc/tc_fwd_sym.cu): a logit of the lower triangle is taken from its mirror tile, where the
    re-rounded k*U operand is the OTHER row -- (k u_a)~ . u_b and (k u_b)~ . u_a differ in the last
    bf16 bit of the operand.  None: decide as the library does."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        M, D = z_t.shape
        N = M // 2
        k = math.log2(math.e) / tau
        flags = capi.FLAG_MIXED_PRECISION if z_t.dtype != torch.bfloat16 else 0
        ub, inv_norm = capi.normalize(z_t, flags)       # the device's own bf16 U and 1/|z|
        uf = ub.float()
        ka = (uf * k).bfloat16().float()                # the A operand: k*U_a, rounded again
        if sym_forward is None:
            sym_forward = M >= 4096 and block % 128 == 0
        idx = torch.arange(M, device=z_t.device)
        pos = (idx + N) % M
        tile = idx // 128
        l2 = torch.empty(M, dtype=torch.float32, device=z_t.device)
        pos_t = torch.empty(M, dtype=torch.float32, device=z_t.device)
        for r0 in range(0, M, block):
            r1 = min(M, r0 + block)
            t = ka[r0:r1] @ uf.t()
            if sym_forward:  # lower-triangle tiles: the mirror tile's value, u_a . (k u_b)~
                lower = tile[r0:r1, None] > tile[None, :]
                t = torch.where(lower, uf[r0:r1] @ ka.t(), t)
                del lower
            if sym_forward:  # captured from the tile's own MMA
                pos_t[r0:r1] = t[idx[: r1 - r0], pos[r0:r1]]
            else:            # a separate fp32 dot product of the operand rows, scaled afterwards
                pos_t[r0:r1] = (uf[r0:r1] * uf[pos[r0:r1]]).sum(dim=1) * k
            t[idx[: r1 - r0], idx[r0:r1]] = float("-inf")
            l2[r0:r1] = torch.log2(torch.exp2(t).double().sum(dim=1)).float()
        ln2 = math.log(2.0)
        lse = l2.double() * ln2
        loss = float(((l2.double() - pos_t.double()) * ln2).mean().item())
        e = torch.exp2(-l2)                              # 2^-l per row / per column
        dz = torch.empty(M, D, dtype=torch.float32, device=z_t.device)
        scale = 1.0 / (tau * M)
        zf = z_t.float()
        for r0 in range(0, M, block):
            r1 = min(M, r0 + block)
            t = ka[r0:r1] @ uf.t()
            w = torch.exp2(t) * (e[r0:r1, None] + e[None, :])
            w[idx[: r1 - r0], idx[r0:r1]] = 0.0
            acc = w.bfloat16().float() @ uf               # the weights are bf16 in TMEM
            du = (acc - 2.0 * uf[pos[r0:r1]]) * scale
            u = zf[r0:r1] * inv_norm[r0:r1, None]         # exact fp32 direction for the Jacobian
            dot = (du * u).sum(dim=1, keepdim=True)
            dz[r0:r1] = grad_out * inv_norm[r0:r1, None] * (du - dot * u)
        return loss, lse, dz
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
