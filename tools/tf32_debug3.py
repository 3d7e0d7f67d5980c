import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ntxent_b200
from ntxent_b200 import capi
torch.manual_seed(0)
M, D, tau = 1024, 128, 0.5
u = capi.normalize(torch.randn(M, D, device="cuda"))[0]
lse_all = capi.forward_rows(u, 0, M, tau, flags=capi.FLAG_FORCE_SIMT)[0]
scale = 1.0 / (tau * M)
pos = (torch.arange(M, device="cuda") + M // 2) % M
def raw(flags):
    du = capi.backward_rows(u, lse_all, 0, M, tau, capi.BWD_CONTRACT, flags=flags)[0]   # row term: pc = 1
    return du / scale + u[pos]
a, b = raw(0), raw(capi.FLAG_FORCE_SIMT)
print("acc simt |max|", b.abs().max().item(), " tc |max|", a.abs().max().item())
print("rel err of acc:", ((a - b).abs().max() / b.abs().max()).item())
print("corr:", torch.corrcoef(torch.stack([a.flatten(), b.flatten()]))[0, 1].item())
# exact fp64 reference of the row term
ud = u.double(); S = ud @ ud.T / tau; S.fill_diagonal_(-1e30)
P = torch.exp(S - lse_all.double()[:, None]); ref = P @ ud
print("simt vs fp64:", ((b.double() - ref).abs().max() / ref.abs().max()).item(), " tc vs fp64:", ((a.double() - ref).abs().max() / ref.abs().max()).item())
# hypotheses: which product does tc compute?  P with tf32-truncated operands?
def trunc(x): return (x.float().view(torch.int32) & ~0x1fff).view(torch.float32).double()
ref_t = trunc(P.float()) @ trunc(u)
print("tc vs fp64 with truncated operands:", ((a.double() - ref_t).abs().max() / ref.abs().max()).item())
for name, alt in (("P^T u", P.T @ ud), ("P u[:, perm?] first 8 K only", P[:, :8] @ ud[:8])):
    print(name, ((a.double() - alt).abs().max() / ref.abs().max()).item())
# per-K-block contributions: does tc miss some K rows?  compare against sum over subsets of 8-row groups
E = (a.double() - ref)
print("error vs ref: per row-tile max", [round(E[t*128:(t+1)*128].abs().max().item() / ref.abs().max().item(), 4) for t in range(8)])
