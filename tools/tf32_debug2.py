import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ntxent_b200
from ntxent_b200 import capi
torch.manual_seed(0)
M, D, tau = 1024, 128, 0.5
u = capi.normalize(torch.randn(M, D, device="cuda"))[0]
lse_all = capi.forward_rows(u, 0, M, tau, flags=capi.FLAG_FORCE_SIMT)[0]
for mode, name in ((capi.BWD_SYMMETRIC, "sym"), (capi.BWD_CONTRACT, "contract")):
    a = capi.backward_rows(u, lse_all, 0, M, tau, mode)
    b = capi.backward_rows(u, lse_all, 0, M, tau, mode, flags=capi.FLAG_FORCE_SIMT)
    for k, nm in ((0, "rows"), (1, "cols")):
        if a[k] is None: continue
        err = (a[k] - b[k]).abs()
        sc = b[k].abs().max()
        print(name, nm, "max rel", (err.max() / sc).item())
        print("  per 32-col group of D:", [round((err[:, g*32:(g+1)*32].max() / sc).item(), 5) for g in range(D // 32)])
        print("  per 128-row tile     :", [round((err[t*128:(t+1)*128].max() / sc).item(), 5) for t in range(M // 128)])
        r = (a[k] / b[k])
        print("  median ratio tc/simt :", r[b[k].abs() > 0.1 * sc].median().item())
