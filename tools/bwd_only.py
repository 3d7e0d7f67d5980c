"""Runs the C4 backward (and forward) a few times; target for ncu captures."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ntxent_b200
from ntxent_b200 import capi
M, D, tau = 32768, 128, 0.1
z = torch.from_numpy(np.random.default_rng(2).standard_normal((M, D)).astype(np.float32)).cuda().to(torch.bfloat16)
ws = capi.workspace(M, D, z, 0)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    capi.forward_backward(z, tau, ws=ws)
torch.cuda.synchronize()
print("done")
