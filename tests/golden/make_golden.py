"""Generates tests/golden/ntxent_golden.npz with PyTorch CPU autograd in fp64.

The reference (/root/reference) cannot produce fixtures: it has no CPU path, does not compile
and its tests hold no golden vectors (SURVEY.md sections 0.1, 0.5, 8c).  These vectors pin the
canonical SimCLR NT-Xent instead -- normalize -> U U^T / tau -> diagonal masked -> cross-entropy
against (i + N) mod M, gradient by autograd -- computed by an implementation that shares no
code with oracle/ or the CUDA kernels.  Inputs use the reference tests' generator shape
(randn, tests/test_utils.hpp:7-14) with fixed seeds, both normalised and raw.

Run:  python tests/golden/make_golden.py
"""
import os

import numpy as np
import torch
import torch.nn.functional as F

CASES = [
    # name, M, D, tau, seed, scale, normalise input first (as the reference tests do)
    ("m4_d2_kat", 4, 2, 1.0, 0, 1.0, False),
    ("m16_d8", 16, 8, 0.5, 1, 1.0, False),
    ("m32_d128_t007_norm", 32, 128, 0.07, 2, 1.0, True),   # tests/test_forward.cpp:14-21 shape
    ("m64_d128_t007_norm", 64, 128, 0.07, 3, 1.0, True),   # 2B rows of B=32
    ("m64_d32_t01", 64, 32, 0.1, 4, 1.0, False),
    ("m128_d64_t05", 128, 64, 0.5, 5, 1.0, False),
    ("m128_d256_t001_big", 128, 256, 0.01, 6, 1e5, True),  # python/test.py:57-79 extremes
    ("m128_d256_t1_small", 128, 256, 1.0, 7, 1e-5, True),
    ("m96_d40_ragged", 96, 40, 0.2, 8, 3.0, False),
]


def ntxent_torch(z, tau):
    M = z.shape[0]
    u = F.normalize(z, dim=1)
    s = u @ u.t() / tau
    s = s.masked_fill(torch.eye(M, dtype=torch.bool), float("-inf"))
    target = (torch.arange(M) + M // 2) % M
    lse = torch.logsumexp(s, dim=1)
    loss = F.cross_entropy(s, target)
    return loss, lse


def main():
    out = {}
    for name, M, D, tau, seed, scale, norm in CASES:
        g = torch.Generator().manual_seed(seed)
        z = torch.randn(M, D, generator=g, dtype=torch.float64)
        if name == "m4_d2_kat":
            z = torch.tensor([[1.0, 0.0], [0.0, 1.0], [1.0, 0.0], [0.0, 1.0]], dtype=torch.float64)
        if norm:
            z = F.normalize(z, dim=1)
        z = (z * scale).detach().requires_grad_(True)
        tau_t = torch.tensor(tau, dtype=torch.float64, requires_grad=True)
        loss, lse = ntxent_torch(z, tau_t)
        dz, dtau = torch.autograd.grad(loss, (z, tau_t))
        out[name + "/z"] = z.detach().numpy()
        out[name + "/tau"] = np.float64(tau)
        out[name + "/loss"] = np.float64(loss.item())
        out[name + "/lse"] = lse.detach().numpy()
        out[name + "/dz"] = dz.numpy()
        out[name + "/dtau"] = np.float64(dtau.item())  # d loss / d tau (learnable temperature)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ntxent_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(CASES), "cases")


if __name__ == "__main__":
    main()
