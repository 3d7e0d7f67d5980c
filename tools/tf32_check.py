import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
import ntxent_b200
from ntxent_b200 import capi
from oracle import oracle as O
from conftest import seeded_z
for (M, D, tau) in [(1030, 64, 0.07), (1030, 128, 0.07), (2048, 64, 0.07), (2048, 64, 0.1), (2048, 128, 0.07)]:
    z = seeded_z(M, D, seed=M + D)
    zt = torch.from_numpy(z).cuda()
    rl, rlse, rdz = O.forward_backward(z.astype(np.float64), tau)
    for name, fl in (("tf32", 0), ("simt", capi.FLAG_FORCE_SIMT)):
        loss, lse = capi.forward(zt, tau, flags=fl)
        dz, _ = capi.backward(zt, lse, None, tau, flags=fl)
        e = np.abs(dz.double().cpu().numpy() - rdz).max() / np.abs(rdz).max()
        print(M, D, tau, name, "loss rel", abs(loss.item() - rl) / rl, "dz err", e)
